"""The N > 1 path on the CPU: two processes over torch.distributed (gloo), each holding one x-slab.

What is checked is the host-side logic the multi-GPU path relies on (csrc/ns_hostgeom.h, compiled into tests/emul for
the host): the slab plan of every rank (column ranges, compact-id ranges, level hierarchy, replicated coarse tail), the
global-index colouring and coefficient tables of a slab, and the halo protocol -- red-black sweeps on two slabs with a
one-column halo exchange before each colour must reproduce the single-slab sweeps bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _cavity(nx, ny):
    import navierstokessolver_b200 as ns
    return ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny))


def _emul(g, rank, world):
    import refapi
    return refapi.Emul(g.hx, g.hy, g.edge_nx, g.edge_ny, g.edge_bc_type, g.edge_bc_info, 0.01, 100.0, rank=rank, world=world)


def _exchange(em, which, rank, world, width):
    """one halo exchange over gloo: send my boundary columns, receive the neighbour's into my halo"""
    for side, peer in ((1, rank + 1), (0, rank - 1)):
        if peer < 0 or peer >= world:
            continue
        mine = torch.from_numpy(em.halo_pack(which, side, width))
        theirs = torch.empty_like(mine)
        if rank < peer:
            dist.send(mine, peer); dist.recv(theirs, peer)
        else:
            dist.recv(theirs, peer); dist.send(mine, peer)
        em.halo_unpack(which, side, width, theirs.numpy())


def _worker(rank, world, port, nx, ny, sweeps, q):
    import refapi
    from refapi import PHI, RHS_PHI
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        g = _cavity(nx, ny)
        em = _emul(g, rank, world)
        info = em.plan_info()
        # every rank sees every plan
        plans = [None] * world
        dist.all_gather_object(plans, info)
        assert [p["i0"] for p in plans] == [r * nx // world for r in range(world)]
        assert [p["i1"] for p in plans] == [(r + 1) * nx // world for r in range(world)]
        assert sum(p["n_local"] for p in plans) == nx * ny
        assert all(plans[r]["id0"] == sum(p["n_local"] for p in plans[:r]) for r in range(world))
        assert len({(p["n_levels"], p["rep_start"]) for p in plans}) == 1
        rep = info["rep_start"]
        for l in range(info["n_levels"]):
            if rep and l >= rep:      # replicated coarse tail: every rank holds the whole level
                assert info["lgi0"][l] == 0 and len({p["lnx"][l] for p in plans}) == 1
            else:                      # distributed: the slabs tile the level and keep >= 6 columns (HX)
                assert [p["lgi0"][l] for p in plans] == [r * plans[0]["lnx"][l] for r in range(world)]
                assert info["lnx"][l] >= 6 or l == 0
        # same seeded global fields on every rank, each takes its id range
        rng = np.random.default_rng(5)
        b = rng.uniform(-1, 1, nx * ny); b -= b.mean()
        x0 = rng.standard_normal(nx * ny)
        lo, hi = info["id0"], info["id0"] + info["n_local"]
        em.set(RHS_PHI, b[lo:hi]); em.set(PHI, x0[lo:hi])
        for _ in range(sweeps):
            for colour in (0, 1):
                _exchange(em, PHI, rank, world, 1)
                em.phi_half_sweep(colour)
        mine = torch.from_numpy(em.get(PHI))
        parts = [torch.empty(p["n_local"], dtype=torch.float64) for p in plans] if rank == 0 else None
        dist.gather(mine, parts, dst=0)
        if rank == 0:
            one = _emul(g, 0, 1)
            one.set(RHS_PHI, b); one.set(PHI, x0)
            one.poisson_sweeps(sweeps)
            got = torch.cat(parts).numpy()
            q.put(("ok", float(np.abs(got - one.get(PHI)).max()), info["n_levels"], rep))
    except Exception as e:   # noqa: BLE001
        if rank == 0:
            q.put(("fail", repr(e), 0, 0))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nx,ny", [(64, 48), (256, 128)])
def test_two_slabs_over_gloo_equal_one_slab(nx, ny):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, nx, ny, 3, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    status, err, n_levels, rep = q.get(timeout=5)
    assert status == "ok", err
    assert err == 0.0                      # bit-identical: same arithmetic, same order
    assert n_levels >= 2
    if nx * ny // 4 <= 4096 * 4:          # small enough that some level reaches the 4096-cell coarse tail
        assert rep >= 1


def test_reference_arm_of_bench_runs_on_cpu(tmp_path):
    """bench.py --impl reference: the reference's own CPU code on a bounded sample, one JSON line with the contract's keys."""
    import json
    import subprocess
    import refapi
    if not refapi.have_ref():
        pytest.skip("oracle/_ref/libnsref.so not built")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-n", "64"],
                         capture_output=True, text=True, timeout=300, cwd=tmp_path)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Gcell-updates/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] == 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0


def test_outflow_ilu_image_is_sliced_per_slab():
    """Several ranks on an outflow domain: every rank factors the GLOBAL ILU(0) of LHS_phi on the host (it is a sequential recurrence
    over all ids) and keeps the slab of w = L U 1 that belongs to its id range -- the slabs of 2 and 4 ranks tile the one-rank w."""
    import refapi
    nx, ny = 96, 24
    hx, hy = np.full(nx, 3.0 / nx), np.linspace(0.8, 1.2, ny) / ny
    args = (hx, hy, [-1, 0, 1, 0], [0, 1, 0, -1], [0, 2, 4, 2], [1.0, 0.0, 0.0, 0.0], 0.01, 100.0)
    whole, ww = refapi.Emul(*args).phi_ilu_w()
    assert ww > 0 and np.isfinite(whole).all()
    for world in (2, 4):
        parts = []
        for rank in range(world):
            em = refapi.Emul(*args, rank=rank, world=world)
            w, ww_r = em.phi_ilu_w()
            assert w.size == nx * ny // world and ww_r == ww
            parts.append(w)
        assert np.array_equal(np.concatenate(parts), whole)
