"""Contract tests on the gloo path, world size 4: rank translation,
non-member no-op + warning, list-on-root-only errors, dtype mismatch, sorted ranks, the
ReduceOp set, async handles."""
import os

import torch

import pdcc_b200 as dist
import _workers

SIZE = 4


def test_contract(tmp_path):
    dist.launch(_workers.contract_checks, size=SIZE, backend="gloo", timeout=240,
                env={"PDCC_TEST_OUT": str(tmp_path)})
    res = [torch.load(os.path.join(tmp_path, f"contract_{r}.pt")) for r in range(SIZE)]
    for r in range(SIZE):
        assert res[r]["all_reduce_SUM"] == 10.0
        assert res[r]["all_reduce_PRODUCT"] == 24.0
        assert res[r]["all_reduce_MAX"] == 4.0
        assert res[r]["all_reduce_MIN"] == 1.0
        assert res[r]["async_sum"] == 6.0 and res[r]["async_done"]
        assert res[r]["avg"] == 1.5 and res[r]["avg_op"] == 1.5
        assert res[r]["scatter_root1"] == [10.0 * r] * 3
        e = res[r]["errors"]
        assert "scatter_list" in e["scatter"] and "gather_list" in e["gather"]
        assert "different dtypes" in e["dtype"]
        assert "torch.Tensor" in e["type"]
        assert "within the world_size" in e["range"]
    assert (res[0]["reduce_SUM"], res[0]["reduce_PRODUCT"], res[0]["reduce_MAX"],
            res[0]["reduce_MIN"]) == (10.0, 24.0, 4.0, 1.0)
    for r in (1, 3):
        assert res[r]["sub_ranks"] == [1, 3]
        assert res[r]["sub_group_rank"] == (0 if r == 1 else 1)
        assert res[r]["sub_bcast"] == 3.0
        assert res[r]["sub_sum"] == 4.0
    for r in (0, 2):
        assert res[r]["non_member_sentinel"] is True
        assert res[r]["non_member_ret"] is None
        assert res[r]["non_member_warned"]
        assert res[r]["non_member_tensor"] == 123.0
    assert res[0]["half_all_gather"] == [0.0, 1.0] and res[3]["half_all_gather"] == [2.0, 3.0]
    assert res[2]["gather_root2"] == [[float(i)] * 2 for i in range(SIZE)]
