"""World-size-2 gloo test of the multi-GPU host logic: disjoint env shards that reproduce the
single-process numbering, and the one collective of the path (sum of the episode-statistics vector, max of the
device time).  Runs on CPU; the per-env code is the CPU build of the device source (tests/hostsim.py)."""
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


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n_per_rank, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from hostsim import HostSim
    from mujoco_jaco_b200.shard import gather_per_rank, max_over_ranks, reduce_stats, shard_range
    off, n = shard_range(rank, world, n_per_rank)
    sim = HostSim(n, task="picking", seed=11, env_offset=off)
    obs0 = sim.reset()
    a = np.random.default_rng(3).uniform(-1, 1, (world * n_per_rank, 7)).astype(np.float32)[off:off + n]
    obs1, rew, done, succ, wb = sim.step(a, auto_reset=True)
    stats = torch.from_numpy(sim.stats.copy())
    local_steps = float(stats[7])
    reduce_stats(stats)
    tmax = max_over_ranks(10.0 + rank)
    table = gather_per_rank(torch.tensor([10.0 + rank, float(rank)]))      # bench.py's per-rank table
    np.savez(os.path.join(outdir, f"r{rank}.npz"), obs0=obs0, obs1=obs1, rew=rew, stats=stats.numpy(),
             local_steps=local_steps, tmax=tmax, table=np.array(table))
    dist.destroy_process_group()


def test_two_rank_shards_match_single_process(tmp_path):
    from hostsim import HostSim
    world, n = 2, 6
    mp.start_processes(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, start_method="spawn")
    ref = HostSim(world * n, task="picking", seed=11, env_offset=0)
    r_obs0 = ref.reset()
    a = np.random.default_rng(3).uniform(-1, 1, (world * n, 7)).astype(np.float32)
    r_obs1, r_rew, *_ = ref.step(a, auto_reset=True)
    for r in range(world):
        z = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(z["obs0"], r_obs0[r * n:(r + 1) * n])      # same Philox streams per global env id
        assert np.array_equal(z["obs1"], r_obs1[r * n:(r + 1) * n])
        assert np.array_equal(z["rew"], r_rew[r * n:(r + 1) * n])
        assert z["local_steps"] == n
        assert np.allclose(z["stats"], ref.stats)                         # all-reduced statistics == global run
        assert z["tmax"] == 11.0
        assert np.array_equal(z["table"], [[10.0, 0.0], [11.0, 1.0]])


def test_shard_range_validation():
    from mujoco_jaco_b200.shard import shard_range
    assert shard_range(3, 8, 65536) == (3 * 65536, 65536)
    with pytest.raises(ValueError):
        shard_range(2, 2, 4)
