"""GPU: checkpoints of a slab run — every rank checkpoints its own local model (parent arrays with the neighbours' halo rows) and a
new run restored from those files continues bit for bit.  Written after the GPU budget was spent; verified with tests/emu (one
process per rank on host cores)."""
import os
import socket

import numpy as np
import pytest

from test_gpu_distributed import _ics, _init_rank, _ranks_available

pytestmark = pytest.mark.gpu


def _run(rank, world, tmp, timestepper, p2p, out):
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    size, FT, dt = (32, 16 * world, 16), np.float64, 2e-3

    def make():
        arch = D.Distributed(D.Partition(y=world))
        return D.DistributedNonhydrostaticModel(arch, FT, size=size, x=(0, 1), y=(0, 2), z=(0, 1), halo=(3, 3, 3),
                                                advection=oc.WENO(FT, 5), timestepper=timestepper, device=rank, p2p=p2p)

    true_model = make()
    true_model.set(**_ics(size, FT))
    true_model.time_step(dt, n_steps=5)
    c = oc.Checkpointer(true_model.model, dir=str(tmp), prefix=f"slab_rank{rank}", overwrite_existing=True)
    path = oc.write_output(c, true_model.model)
    true_model.time_step(dt, n_steps=4)
    restarted = make()
    oc.set_from_checkpoint(restarted.model, path)          # set!(model, filepath); no update_state!: the halo rows are in the file
    restarted.time_step(dt, n_steps=4)
    assert restarted.clock.iteration == true_model.clock.iteration == 9 and restarted.clock.time == true_model.clock.time
    for n in ("u", "v", "w", "pNHS"):
        a, b = restarted.gather(n), true_model.gather(n)
        if rank == 0:
            out[n] = bool(np.array_equal(a, b))


def _worker(rank, world, port, tmp, timestepper, p2p, path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    _init_rank(rank, world)
    try:
        out = {}
        _run(rank, world, tmp, timestepper, p2p, out)
        if rank == 0:
            np.savez(path, **out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_one_rank_slab_run_continues_from_its_checkpoint(timestepper, p2p, tmp_path):
    out = {}
    _run(0, 1, tmp_path, timestepper, p2p, out)
    assert out and all(out.values()), out


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_slab_run_continues_from_per_rank_checkpoints(world, timestepper, tmp_path):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    path = str(tmp_path / "ok.npz")
    mp.spawn(_worker, args=(world, port, str(tmp_path), timestepper, True, path), nprocs=world, join=True)
    res = dict(np.load(path))
    assert res and all(bool(v) for v in res.values()), res
