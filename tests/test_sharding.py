"""Multi-GPU sharding on CPU: world_size-2 gloo processes drive the host logic
over the numpy oracle; the same code runs with NCCL on the GPUs (bench.py --gpus N)."""
import socket

import numpy as np
import pytest

from cyten_b200.sharding import lpt_assign, lpt_imbalance


def test_lpt_assign_is_deterministic_and_balanced():
    costs = np.asarray([520, 487, 487, 421, 421, 330, 330, 234, 234, 151, 151, 88, 88, 46, 46, 22, 22, 9, 9], float) ** 3
    for parts in (1, 2, 4, 8):
        owner = lpt_assign(costs, parts)
        assert np.array_equal(owner, lpt_assign(costs, parts))
        assert set(owner.tolist()) <= set(range(parts))
        # greedy LPT is within 4/3 of optimal; the optimum is at least max(mean load, largest item)
        load = np.bincount(owner, weights=costs, minlength=parts)
        assert load.max() <= 4 / 3 * max(costs.sum() / parts, costs.max()) + 1e-9
    assert lpt_imbalance(costs, lpt_assign(costs, 1), 1) == 1.0
    # ties: lower index first, to the lower part id
    assert lpt_assign([1, 1, 1, 1], 2).tolist() == [0, 1, 0, 1]
    assert lpt_assign([], 3).tolist() == []


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_ops_match_single_process(tmp_path, world):
    import torch.multiprocessing as mp
    from _sharding_worker import worker
    mp.spawn(worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert len(list(tmp_path.glob("rank*.npy"))) == world


@pytest.mark.gpu
@pytest.mark.parametrize("chi", [64, 1024])
def test_sharded_ops_match_unsharded_on_gpus_nccl(tmp_path, chi):
    """The same checks with one rank per real GPU over NCCL and the CUDA library (column-split matvec, sharded tdot,
    sharded truncated SVD against the unsharded result of the same GPU).  Needs >= 2 GPUs (gpurun --gpus 2)."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    from _sharding_worker import worker
    world = 2
    mp.spawn(worker, args=(world, _free_port(), str(tmp_path), "nccl", chi), nprocs=world, join=True)
    assert len(list(tmp_path.glob("rank*.npy"))) == world
