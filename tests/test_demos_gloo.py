"""CPU plumbing: the reference's seven demos through
init_process(..., backend="gloo") at world size 4, checked against README.md's output."""
import os

import pytest
import torch

import pdcc_b200 as dist
import _workers

SIZE = 4


def _run(name, tmp_path, capfd, monkeypatch):
    monkeypatch.setenv("PDCC_TEST_OUT", str(tmp_path))
    monkeypatch.setenv("PDCC_TEST_DEMO", name)
    codes = dist.launch(_workers.run_demo, size=SIZE, backend="gloo", timeout=180,
                        env={"PDCC_TEST_OUT": str(tmp_path), "PDCC_TEST_DEMO": name})
    assert list(codes) == [0] * SIZE
    out = capfd.readouterr().out
    res = [torch.load(os.path.join(tmp_path, f"{name}_{r}.pt")) for r in range(SIZE)]
    return out, res


def test_hello_world(tmp_path, capfd, monkeypatch):
    out, _ = _run("hello_world", tmp_path, capfd, monkeypatch)
    for r in range(SIZE):
        assert f"[{r}] say hi!" in out  # README.md:76-81


def test_do_reduce(tmp_path, capfd, monkeypatch):
    out, res = _run("do_reduce", tmp_path, capfd, monkeypatch)
    assert res[0].item() == 4.0  # README.md:105-112: only rank 0 is a contract
    assert "[0] data = 4.0" in out


def test_do_all_reduce(tmp_path, capfd, monkeypatch):
    out, res = _run("do_all_reduce", tmp_path, capfd, monkeypatch)
    for r in range(SIZE):
        assert res[r].item() == 4.0
        assert f"[{r}] data = 4.0" in out  # README.md:140-145


def test_do_scatter(tmp_path, capfd, monkeypatch):
    out, res = _run("do_scatter", tmp_path, capfd, monkeypatch)
    for r in range(SIZE):
        assert res[r].item() == r + 1
        assert f"[{r}] data = {float(r + 1)}" in out  # README.md:175-180


def test_do_gather(tmp_path, capfd, monkeypatch):
    out, res = _run("do_gather", tmp_path, capfd, monkeypatch)
    assert [t.item() for t in res[0]] == [0.0, 1.0, 2.0, 3.0]
    assert "[0] data = [tensor([0.]), tensor([1.]), tensor([2.]), tensor([3.])]" in out
    assert "[1] data" not in out  # README.md:211-213: only rank 0 prints


def test_do_all_gather(tmp_path, capfd, monkeypatch):
    out, res = _run("do_all_gather", tmp_path, capfd, monkeypatch)
    for r in range(SIZE):
        assert [t.item() for t in res[r]] == [0.0, 1.0, 2.0, 3.0]
        assert f"[{r}] data = [tensor([0.]), tensor([1.]), tensor([2.]), tensor([3.])]" in out


def test_do_broadcast(tmp_path, capfd, monkeypatch):
    out, res = _run("do_broadcast", tmp_path, capfd, monkeypatch)
    for r in range(SIZE):
        assert res[r].item() == 0.0
        assert f"[{r}] data = tensor([0.])" in out  # README.md:279-284


def test_launcher_reports_failed_rank():
    with pytest.raises(RuntimeError, match="rank 1 failed with exit code 3"):
        dist.launch(_workers.crash_rank1, size=2, backend="gloo", timeout=120)
