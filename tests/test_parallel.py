"""N > 1 path on CPU: world_size-2 gloo processes, each integrating half the batch with the torch-CPU emulation of the
backend (tests/fake_backend.py).  Checks SURVEY.md 8e: with the error-norm partial sums all-reduced (SUM) both ranks
take exactly the step sequence of one process on the global batch (the reference's batch-shared dopri5 step size),
and the flat gradient bucket averages to the single-process gradient."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _build(scale):
    import ode_rl_b200 as odex
    from tests.util import make_ode_func

    torch.manual_seed(0)
    f = make_ode_func(odex, C=32, n_units=32, scale=scale)
    return odex, f


def _worker(rank, world, port, mode, q):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ode_rl_b200 import kernels, parallel
        from tests.fake_backend import FakeBackend
        from tests.util import timesteps

        kernels.set_backend(FakeBackend())
        odex, f = _build(3.0)
        _, tp = timesteps(10, 10)
        torch.manual_seed(5)
        z_all = torch.randn(5, 32, 8, 16)                    # ragged: 3 + 2 samples
        z_all[3:] *= 2.0                                     # make the shards' error norms differ
        lo, hi = parallel.shard_range(5, rank, world)
        # ---- dopri5 forward with the shared step size
        solver = parallel.shard_solver(odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4), mode=mode)
        assert solver.reduce_fn is not None
        with torch.no_grad():
            sol = solver(z_all[lo:hi], tp)
        st = solver.last_stats
        # ---- rk4 training gradient through the flat bucket
        rk = odex.DiffEqSolver(f, "rk4")
        z = z_all[lo:hi].clone()
        # staged bucket: the first half of the parameters is reduced from a post-accumulate hook during the backward, the rest
        # by allreduce() -- same averaged gradients as the single flat bucket
        plist = list(rk.parameters())
        bucket = parallel.GradBucket(plist, stages=[plist[:len(plist) // 2]])
        out = rk(z, tp[:3])
        (out.square().sum() / 5.0).backward()                # global mean over 5 samples -> per-rank partial sums
        assert any(bucket._launched)
        nbytes = bucket.allreduce()
        assert not any(bucket._launched)
        g = torch.cat([p.grad.reshape(-1) for p in rk.parameters()]) * world   # bucket averages; undo for the sum
        q.put((rank, sol.contiguous().numpy().copy(), [(a[1], a[3]) for a in st.attempts], (st.n_accept, st.n_reject),
               g.detach().numpy().copy(), nbytes))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["sum", "max"])
def test_two_rank_gloo_shared_steps_and_grad_bucket(mode):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        r = q.get(timeout=240)
        res[r[0]] = (r[0], torch.from_numpy(r[1]), r[2], r[3], torch.from_numpy(r[4]), r[5])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0

    # both ranks took the same attempts
    assert res[0][2] == res[1][2] and res[0][3] == res[1][3]

    # single process on the global batch
    from ode_rl_b200 import kernels
    from tests.fake_backend import FakeBackend
    from tests.util import timesteps

    prev = kernels.set_backend(FakeBackend())
    try:
        odex, f = _build(3.0)
        _, tp = timesteps(10, 10)
        torch.manual_seed(5)
        z_all = torch.randn(5, 32, 8, 16)
        z_all[3:] *= 2.0
        solver = odex.DiffEqSolver(f, "dopri5", odeint_rtol=1e-3, odeint_atol=1e-4)
        with torch.no_grad():
            sol = solver(z_all, tp)
        st = solver.last_stats
        rk = odex.DiffEqSolver(f, "rk4")
        out = rk(z_all.clone(), tp[:3])
        (out.square().sum() / 5.0).backward()
        g = torch.cat([p.grad.reshape(-1) for p in rk.parameters()])
    finally:
        kernels.set_backend(prev)

    sharded = torch.cat([res[0][1], res[1][1]], dim=1)
    if mode == "sum":
        # exactly the global-batch sequence: same accept/reject pattern, dt equal to fp64 summation-order noise
        assert res[0][3] == (st.n_accept, st.n_reject)
        for (dt_a, acc_a), b in zip(res[0][2], st.attempts):
            assert bool(acc_a) == bool(b[3]) and abs(dt_a - b[1]) <= 1e-9 * abs(b[1])
        assert (sharded - sol).abs().max() <= 1e-5 * sol.abs().max()
    else:
        # max mode: rank-consistent, at least as conservative (never fewer attempts), still within tolerance
        assert sum(res[0][3]) >= st.n_accept + st.n_reject
        assert (sharded - sol).abs().max() <= 5e-3 * sol.abs().max()
    assert res[0][5] == g.numel() * 4
    # fp32 summation order differs between (3 + 2 samples, summed by the bucket) and one pass over 5 samples
    atol = 1e-5 * float(g.abs().max())
    assert torch.allclose(res[0][4], g, rtol=1e-4, atol=atol) and torch.allclose(res[1][4], g, rtol=1e-4, atol=atol)


def test_shard_range_ragged_and_empty():
    from ode_rl_b200.parallel import shard_range

    assert [shard_range(5, r, 2) for r in range(2)] == [(0, 3), (3, 5)]
    assert [shard_range(2, r, 4) for r in range(4)] == [(0, 1), (1, 2), (2, 2), (2, 2)]
    assert shard_range(0, 0, 1) == (0, 0)
    covered = []
    for r in range(8):
        lo, hi = shard_range(513, r, 8)
        covered += list(range(lo, hi))
    assert covered == list(range(513))
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


# ---- BASELINE configs[4]: VidODE sharded by batch (bench.py --workload vidode --gpus N) -------------------------------------
def _vidode_model_and_batch():
    import numpy as np
    from ode_rl_b200.harness import VidODEHarness

    torch.manual_seed(0)
    model = VidODEHarness("dopri5", in_channels=3, resolution=32).eval()      # eval: BatchNorm per sample, so shards == whole
    with torch.no_grad():
        for p in model.ode_decoder_func.parameters():
            p.mul_(4.0)                                                       # several attempts, some rejected
    B, T = 4, 3
    frames = torch.rand(B, T, 3, 32, 32, generator=torch.Generator().manual_seed(7))
    frames[2:] *= 3.0                                                         # the shards' error norms differ
    ts = torch.tensor(np.arange(0, 2 * T) / (2 * T))
    batch = {"observed_tp": ts[:T], "tp_to_predict": ts[T:], "observed_mask": torch.ones(B, T, 1),
             "mask_predicted_data": torch.ones(B, T, 1)}
    return model, frames, batch


def _vidode_worker(rank, world, port, q):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ode_rl_b200 import kernels, parallel
        from tests.fake_backend import FakeBackend

        kernels.set_backend(FakeBackend())
        model, frames, batch = _vidode_model_and_batch()
        parallel.shard_solver(model.diffeq_solver, mode="sum")
        lo, hi = parallel.shard_range(frames.shape[0], rank, world)
        with torch.no_grad():
            pred, extra = model(frames[lo:hi], batch)
        st = model.diffeq_solver.last_stats
        q.put((rank, pred.contiguous().numpy().copy(), extra["sol_y"].contiguous().numpy().copy(),
               [(a[1], a[3]) for a in st.attempts]))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_vidode_forward_equals_single_process():
    """Each rank runs the VidODE forward on its half of the sequences with the dopri5 error norm all-reduced (SUM): the
    ranks take the step sequence of one process on the whole batch and the concatenated predictions equal its output."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_vidode_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        r = q.get(timeout=240)
        res[r[0]] = r
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res[0][3] == res[1][3]

    from ode_rl_b200 import kernels
    from tests.fake_backend import FakeBackend

    prev = kernels.set_backend(FakeBackend())
    try:
        model, frames, batch = _vidode_model_and_batch()
        with torch.no_grad():
            pred, extra = model(frames, batch)
        st = model.diffeq_solver.last_stats
    finally:
        kernels.set_backend(prev)
    assert len(st.attempts) == len(res[0][3]) and st.n_reject + st.n_accept >= 2
    for (dt_a, acc_a), b in zip(res[0][3], st.attempts):
        assert bool(acc_a) == bool(b[3]) and abs(dt_a - b[1]) <= 1e-9 * abs(b[1])
    sol = torch.cat([torch.from_numpy(res[0][2]), torch.from_numpy(res[1][2])], dim=0)
    prd = torch.cat([torch.from_numpy(res[0][1]), torch.from_numpy(res[1][1])], dim=0)
    assert (sol - extra["sol_y"]).abs().max() <= 1e-5 * extra["sol_y"].abs().max()
    assert (prd - pred).abs().max() <= 1e-5 * pred.abs().max()
