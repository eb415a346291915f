"""The N>1 host logic on CPU: world_size-2 and -4 `gloo` jobs that run the solver's multi-rank SCHEME with the
oracle's operators in place of the CUDA kernels:

  * block ranges from the product library's pb200_decompose_2d / pb200_choose_process_grid,
  * neighbours by the reference's rule (S4:338-344: px = rank % Px, py = rank / Px),
  * only z = D^-1 r crosses rank boundaries (one halo exchange per iteration); p on the ring is recomputed
    locally as z + beta*p, exactly what k1_kernel does on the GPU,
  * the three scalars are all-reduced (gloo standing in for NCCL).

It must reproduce the reference's multi-rank iteration counts (tests/golden) and the single-rank solution.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def allreduce(x):
    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t)
    return float(t[0])


def exchange_z(z, nbr):
    """First/last interior rows and columns of z go to the neighbours' ring (rows lj=1..ny, columns li=1..nx)."""
    nx, ny = z.shape[0] - 2, z.shape[1] - 2
    plan = [("left", z[1, 1:ny + 1], (0, slice(1, ny + 1))), ("right", z[nx, 1:ny + 1], (nx + 1, slice(1, ny + 1))),
            ("down", z[1:nx + 1, 1], (slice(1, nx + 1), 0)), ("up", z[1:nx + 1, ny], (slice(1, nx + 1), ny + 1))]
    reqs, recvs = [], []
    for name, send, where in plan:
        peer = nbr[name]
        if peer < 0:
            continue
        sbuf = torch.from_numpy(np.ascontiguousarray(send))
        rbuf = torch.empty_like(sbuf)
        reqs.append(dist.isend(sbuf, peer))
        reqs.append(dist.irecv(rbuf, peer))
        recvs.append((rbuf, where))
    for r in reqs:
        r.wait()
    for rbuf, where in recvs:
        z[where] = rbuf.numpy()


def worker(rank, world, port, M, N, Px, Py, out_dir):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from __graft_entry__ import load_package
    from oracle_lib import Oracle
    pb, orc = load_package(), Oracle()
    if Px <= 0:
        Px, Py = pb.choose_process_grid(world)
    i0, i1, j0, j1 = pb.decompose_2d(M, N, Px, Py, rank)
    px, py = rank % Px, rank // Px
    nbr = {"left": rank - 1 if px > 0 else -1, "right": rank + 1 if px < Px - 1 else -1,
           "down": rank - Px if py > 0 else -1, "up": rank + Px if py < Py - 1 else -1}
    h1, h2 = 2.0 / M, 1.2 / N
    a, b, B = orc.setup(M, N, (i0, i1, j0, j1))
    w, r = np.zeros_like(B), B.copy()
    inner = (slice(1, -1), slice(1, -1))
    z = np.zeros_like(B)
    z[inner] = orc.apply_Dinv(r, a, b, h1, h2)[inner]
    exchange_z(z, nbr)
    p = np.zeros_like(B)
    zr_old = allreduce(orc.dot(z, r, h1, h2))
    iters, converged = 0, False
    for k in range(1, (M - 1) * (N - 1) + 1):
        iters = k
        p = z.copy() if k == 1 else z + beta * p          # ring included: no exchange of p
        Ap = orc.apply_A(p, a, b, h1, h2)
        denom = allreduce(orc.dot(Ap, p, h1, h2))
        if abs(denom) < 1e-15:
            break
        alpha = zr_old / denom
        diff2 = orc.update_w_r(w, r, p, Ap, alpha)
        z[inner] = orc.apply_Dinv(r, a, b, h1, h2)[inner]
        exchange_z(z, nbr)
        zr_new = allreduce(orc.dot(z, r, h1, h2))
        diff = np.sqrt(allreduce(diff2) * h1 * h2)
        if diff < 1e-6:
            converged = True
            break
        beta = zr_new / zr_old
        zr_old = zr_new
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), w=w[inner], blk=np.array([i0, i1, j0, j1]), iters=iters,
             converged=converged)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("M,N,world,Px,Py,key", [(40, 40, 2, 0, 0, "40x40@2"), (40, 40, 2, 2, 1, "40x40@2"),
                                                 (37, 53, 4, 0, 0, None), (40, 40, 4, 0, 0, "40x40@4")])
def test_gloo_multirank_scheme(tmp_path, golden, oracle, M, N, world, Px, Py, key):
    port = free_port()
    mp.spawn(worker, args=(world, port, M, N, Px, Py, str(tmp_path)), nprocs=world, join=True)
    res1, w1, _ = oracle.solve(M, N, want_w=True)
    w = np.zeros((M + 1, N + 1))
    for rank in range(world):
        d = np.load(tmp_path / f"rank{rank}.npz")
        i0, i1, j0, j1 = d["blk"]
        w[i0:i1 + 1, j0:j1 + 1] = d["w"]
        assert int(d["iters"]) == res1.iterations and bool(d["converged"])
    if key:
        assert res1.iterations == golden["iters_stage3_ranks"][key]     # the reference on the same rank count
    np.testing.assert_allclose(w, w1, rtol=0, atol=1e-12)
