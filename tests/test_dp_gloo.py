"""D1 on the CPU plumbing path: parameter broadcast + bucketed fused gradient averaging give
the same parameters as one process training on the concatenated batch."""
import os

import torch

import pdcc_b200 as dist
import _workers


def test_data_parallel_matches_big_batch(tmp_path):
    dist.launch(_workers.dp_mlp_check, size=2, backend="gloo", timeout=240,
                env={"PDCC_TEST_OUT": str(tmp_path)})
    res = [torch.load(os.path.join(tmp_path, f"dp_{r}.pt")) for r in range(2)]
    for r in res:
        assert r["buckets"] >= 3
        assert r["err"] < 1e-5, r
