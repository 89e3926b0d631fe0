"""API-layer bookkeeping (ranks, sizes, group translation, lifecycle errors) on gloo."""
import os

import pytest
import torch

import pdcc_b200 as dist
import _workers


def test_api_misc(tmp_path):
    n = 3
    dist.launch(_workers.api_misc, size=n, backend="gloo", timeout=240,
                env={"PDCC_TEST_OUT": str(tmp_path)})
    res = [torch.load(os.path.join(tmp_path, f"api_{r}.pt")) for r in range(n)]
    for r in range(n):
        assert (res[r]["rank"], res[r]["world"], res[r]["backend"]) == (r, n, "gloo")
        assert res[r]["initialized"] and "twice" in res[r]["double_init"]
        assert "unsupported reduce op" in res[r]["bad_op"]
        assert "Invalid process group" in res[r]["bad_group"]
        assert "init_process_group('b200'" in res[r]["native_on_gloo"]
        assert res[r]["sym_fallback"] == ((3, 5), "torch.float64", False)
    for r in (0, n - 1):
        assert res[r]["group_size"] == 2 and res[r]["group_rank"] == (0 if r == 0 else 1)
        assert res[r]["global_of_1"] == n - 1 and res[r]["group_of_last"] == 1
        assert "not part of group" in res[r]["bad_src"]
        assert res[r]["bcast_group_src"] == float(n - 1) and res[r]["barrier_done"]
    assert res[0]["reduce_group_dst"] == 2.0
    assert res[1]["sentinel_rank"] == -1 and res[1]["sentinel_size"] == -1


def test_uninitialised_and_bad_backends():
    assert not dist.is_initialized()
    with pytest.raises(ValueError, match="has not been initialized"):
        dist.get_rank()
    with pytest.raises(ValueError, match="NCCL"):
        dist.init_process_group("nccl", rank=0, world_size=1)
    with pytest.raises(ValueError, match="unknown backend"):
        dist.init_process_group("mpi", rank=0, world_size=1)
    assert not dist.is_initialized()
