"""The N>1 path on CPU: two gloo processes, host test double for the kernels.

Exercises the product's multi-rank host logic -- index-level batch sharding with
even_batches, bucketed gradient all-reduce from autograd hooks with no_sync semantics on
accumulation steps, 1/W folded into the step, scheduler stepped W times, loss mean over
ranks, Meter counts with padding masked -- and compares with golden vectors produced by the
UNMODIFIED reference under real torch DDP over gloo (tests/golden/make_golden_ws2.py).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import pytest
import torch.multiprocessing as mp

import dist_worker

GOLDEN = Path(__file__).parent / "golden"
_PORT = [29710]


def _run(case: str, tmp_path, device_metric: bool):
    _PORT[0] += 1
    mp.spawn(dist_worker.engine_worker, args=(2, _PORT[0], case, str(tmp_path), device_metric),
             nprocs=2, join=True)
    return [np.load(tmp_path / f"{case}_rank{r}.npz") for r in range(2)]


@pytest.mark.parametrize("case", list(dist_worker.WS2_CASES))
@pytest.mark.parametrize("device_metric", [False, True])
def test_two_ranks_match_reference_under_ddp(case, device_metric, tmp_path):
    gold = np.load(GOLDEN / f"{case}.npz")
    r0, r1 = _run(case, tmp_path, device_metric)
    # replicas stay bit-identical: same gradients after the exchange, deterministic norm
    for k in gold.files:
        if k.startswith("param/"):
            assert np.array_equal(r0[k].view(np.int32), r1[k].view(np.int32)), k
    assert r0["loss_steps"].tolist() == gold["loss_steps"].tolist()
    assert r0["lr_values"].tolist() == gold["lr_values"].tolist()      # scheduler stepped W x
    np.testing.assert_allclose(r0["loss_values"], gold["loss_values"], rtol=1e-6)
    assert r0["accuracy_history"].tolist() == gold["accuracy_history"].tolist()
    assert r1["accuracy_history"].tolist() == gold["accuracy_history"].tolist()
    for k in gold.files:
        if k.startswith("param/"):
            want = gold[k]
            assert np.abs(r0[k] - want).max() <= 2e-5 * np.abs(want).max(), k


def test_two_ranks_fp16_skip_the_same_steps_as_reference_under_ddp(tmp_path):
    """mixed_precision="fp16" on two ranks: the golden is the unmodified reference under DDP
    with one torch GradScaler per rank; its first steps overflow.  Both engine ranks must skip
    exactly those steps (lr series; the scheduler moves W times per real step) and end on the
    reference's parameters bit for bit (SGD: no sqrt between the oracle and torch)."""
    case = "ws2_fp16_sgd"
    gold = np.load(GOLDEN / f"{case}.npz")
    r0, r1 = _run(case, tmp_path, True)
    assert len(set(gold["lr_values"][:5].tolist())) == 1, "the golden should start with skips"
    assert r0["lr_values"].tolist() == gold["lr_values"].tolist()
    assert r0["loss_steps"].tolist() == gold["loss_steps"].tolist()
    np.testing.assert_allclose(r0["loss_values"], gold["loss_values"], rtol=1e-6)
    assert r0["accuracy_history"].tolist() == gold["accuracy_history"].tolist()
    for k in gold.files:
        if k.startswith("param/"):
            assert np.array_equal(r0[k].view(np.int32), r1[k].view(np.int32)), k
            assert np.array_equal(r0[k].view(np.int32), gold[k].view(np.int32)), k


def test_sharded_batch_sampler_even_batches():
    from rocket_b200.staging import ShardedBatchSampler
    from torch.utils.data import BatchSampler, SequentialSampler

    def shards(n, bs, world):
        base = BatchSampler(SequentialSampler(range(n)), batch_size=bs, drop_last=False)
        return [list(ShardedBatchSampler(base, world, r)) for r in range(world)]

    assert shards(10, 2, 2) == [[[0, 1], [4, 5], [8, 9]], [[2, 3], [6, 7], [0, 1]]]
    assert shards(9, 2, 2) == [[[0, 1], [4, 5], [8, 0]], [[2, 3], [6, 7], [1, 2]]]
    assert shards(8, 2, 2) == [[[0, 1], [4, 5]], [[2, 3], [6, 7]]]
    for n, bs, w in ((333, 64, 2), (7, 3, 4), (64, 8, 8), (1, 4, 2)):
        out = shards(n, bs, w)
        assert len({len(s) for s in out}) == 1                 # same number of batches
        assert all(len(b) == bs for s in out for b in s)       # all full sized
        seen = [i for step in zip(*out) for b in step for i in b]
        assert seen[:n] == list(range(n))                      # gather order = dataset order
