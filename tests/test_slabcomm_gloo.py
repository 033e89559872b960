"""CPU, world_size 2, gloo: the communication plumbing of the y-slab driver (ring halo exchange,
all-to-all transposes with the chunk layout of mhh_transpose_copy, max/sum reductions)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from microhh_b200.multigpu import SlabComm
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    comm = SlabComm(rank, world)
    ok = True
    # ring halo: what I send north must arrive as my north neighbour's south ghost
    to_n, to_s = torch.full((6,), 10. + rank, dtype=torch.float64), torch.full((6,), 20. + rank, dtype=torch.float64)
    fs, fn = torch.zeros(6, dtype=torch.float64), torch.zeros(6, dtype=torch.float64)
    comm.halo_exchange(to_n, to_s, fs, fn)
    ok &= bool((fs == 10. + (rank - 1) % world).all()) and bool((fn == 20. + (rank + 1) % world).all())
    # exec_xy then exec_yx with the chunk layout [r][k][j][ii] must round-trip and realise the global transpose
    itot, jtot, kmax = 8, 6, 3
    jmax, iblock = jtot // world, itot // world
    glob = np.arange(kmax * jtot * itot, dtype=np.float64).reshape(kmax, jtot, itot)
    X = glob[:, rank * jmax:(rank + 1) * jmax, :]
    send = np.stack([X[:, :, r * iblock:(r + 1) * iblock] for r in range(world)])            # [r][k][j][ii]
    recv = torch.zeros(send.size, dtype=torch.float64)
    comm.all_to_all(recv, torch.from_numpy(np.ascontiguousarray(send)).reshape(-1))
    chunks = recv.numpy().reshape(world, kmax, jmax, iblock)
    Y = np.concatenate([chunks[r] for r in range(world)], axis=1)                             # [k][jtot][iblock]
    ok &= np.array_equal(Y, glob[:, :, rank * iblock:(rank + 1) * iblock])
    back = torch.zeros(send.size, dtype=torch.float64)
    send2 = np.stack([Y[:, r * jmax:(r + 1) * jmax, :] for r in range(world)])
    comm.all_to_all(back, torch.from_numpy(np.ascontiguousarray(send2)).reshape(-1))
    c2 = back.numpy().reshape(world, kmax, jmax, iblock)
    ok &= np.array_equal(np.concatenate([c2[r] for r in range(world)], axis=2), X)
    # the same transpose cut into k ranges, each with its own chunk buffer [c][r][k' < nk][j][ii] and its own
    # asynchronous all-to-all (SlabPres.kranges): the segments must reassemble the same Y
    nc, nk = 3, 1
    seg = nk * jmax * itot
    sendk = np.concatenate([np.stack([X[c * nk:(c + 1) * nk, :, r * iblock:(r + 1) * iblock] for r in range(world)]).reshape(-1)
                            for c in range(nc)])
    send_t, recv_t = torch.from_numpy(np.ascontiguousarray(sendk)), torch.zeros(sendk.size, dtype=torch.float64)
    works = [comm.all_to_all_async(recv_t[c * seg:(c + 1) * seg], send_t[c * seg:(c + 1) * seg]) for c in range(nc)]
    for w in works:
        if w is not None:
            w.wait()
    Yk = np.concatenate([np.concatenate(list(recv_t.numpy()[c * seg:(c + 1) * seg].reshape(world, nk, jmax, iblock)), axis=1)
                         for c in range(nc)], axis=0)
    ok &= np.array_equal(Yk, Y)
    ok &= comm.allreduce_max(float(rank), "cpu") == world - 1
    ok &= float(comm.allreduce_sum([1.0, float(rank)], "cpu")[1]) == sum(range(world))
    # the pair of dt maxima in one collective; a NaN on ANY rank must come out as NaN on every rank (the CFL guard)
    m = comm.allreduce_max_many([1.5 + rank, 7. - rank], "cpu")
    ok &= m == [1.5 + world - 1, 7.]
    m = comm.allreduce_max_many([float("nan") if rank == 1 else 3., 2.], "cpu")
    ok &= all(v != v for v in m)
    q.put((rank, ok))
    dist.destroy_process_group()


def test_slabcomm_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29641, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(ok for _, ok in res), res
