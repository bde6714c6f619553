"""ProcessGroupManager over the [DP, PP, CP, EP, TP] grid (reference tests: tests/parallel/test_process_group.py)."""
import pytest
import torch

from dist_utils import run_distributed


def _grid_worker(rank, world):
    from scaletorch_b200.parallel.process_group import (get_process_group_manager, process_group_manager as pgm,
                                                        setup_process_group_manager)
    import torch.distributed as dist

    assert not pgm
    setup_process_group_manager(tp_size=2, cp_size=1, pp_size=2, dp_size=2, ep_size=1)
    assert pgm and get_process_group_manager() is not None
    # TP fastest, DP slowest
    assert pgm.tp_rank == rank % 2 and pgm.pp_rank == (rank // 2) % 2 and pgm.dp_rank == rank // 4
    assert pgm.tp_group_ids == [rank - rank % 2, rank - rank % 2 + 1]
    assert pgm.world_size == 8 and pgm.grid.shape == (2, 2, 1, 1, 2)
    assert pgm.pp_is_first_stage == (pgm.pp_rank == 0) and pgm.pp_is_last_stage == (pgm.pp_rank == 1)
    if pgm.pp_is_first_stage:
        assert pgm.pp_next_rank == rank + 2 and pgm.pp_prev_rank is None
    assert pgm.cp_send_rank == rank and pgm.cp_recv_rank == rank  # cp = 1 ring
    assert pgm.cp_dp_world_size == 2 and pgm.pp_dp_world_size == 4 and pgm.grad_sync_world_size == 2
    t = torch.ones(1)
    dist.all_reduce(t, group=pgm.tp_group)
    assert t.item() == 2
    t = torch.tensor([float(rank)])
    dist.all_reduce(t, group=pgm.dp_group)
    assert t.item() == rank % 4 + (rank % 4 + 4)
    info = pgm.get_info()
    assert info["tp_world_size"] == 2 and "TP(2)" in str(pgm)
    pgm.cleanup()


def test_grid_coordinates_and_groups_world8():
    run_distributed(_grid_worker, 8)


def _ep_worker(rank, world):
    from scaletorch_b200.parallel.process_group import process_group_manager as pgm, setup_process_group_manager

    setup_process_group_manager(tp_size=1, cp_size=1, pp_size=1, dp_size=2, ep_size=2)
    assert pgm.ep_rank == rank % 2 and pgm.dp_rank == rank // 2
    assert pgm.data_world_size == 4 and pgm.data_rank == rank          # distinct data per EP rank by default
    assert pgm.grad_sync_world_size == 4 and pgm.expert_grad_sync_world_size == 2


def test_ep_axis_and_data_ranks():
    run_distributed(_ep_worker, 4)


def _bad_worker(rank, world):
    from scaletorch_b200.parallel.process_group import setup_process_group_manager

    with pytest.raises(ValueError, match="World size"):
        setup_process_group_manager(tp_size=2, cp_size=2, pp_size=1, dp_size=1)
    with pytest.raises(ValueError, match="positive"):
        setup_process_group_manager(tp_size=0, cp_size=1, pp_size=1, dp_size=2)


def test_validation_errors():
    run_distributed(_bad_worker, 2)


def test_requires_initialised_dist_and_proxy_is_falsy():
    from scaletorch_b200.parallel.process_group import ProcessGroupManager, process_group_manager as pgm

    assert not pgm
    with pytest.raises(RuntimeError):
        ProcessGroupManager(1, 1, 1, 1)
    with pytest.raises(AttributeError):
        _ = pgm.tp_rank
