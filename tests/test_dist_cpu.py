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


def test_sharded_batch_sampler_known_cases_from_accelerate():
    """Expected shards written out by hand, the cases accelerate's own
    tests/test_data_loader.py::test_batch_sampler_shards_with_no_splits checks (recalled, the
    package is not on the image): batch size 3, two processes, even_batches.  Both the engine's
    sampler and the test shim's restatement (which the goldens were generated through) must
    give them -- the one place where accelerate-level semantics are pinned by something other
    than this repository's own restatement."""
    import sys
    from torch.utils.data import BatchSampler
    import oracle
    from rocket_b200.staging import ShardedBatchSampler
    if str(oracle.SHIMS) not in sys.path:
        sys.path.insert(0, str(oracle.SHIMS))
    from accelerate import BatchSamplerShard          # the shim's class

    cases = {
        24: ([[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
             [[3, 4, 5], [9, 10, 11], [15, 16, 17], [21, 22, 23]]),
        21: ([[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],          # odd number of batches
             [[3, 4, 5], [9, 10, 11], [15, 16, 17], [0, 1, 2]]),
        22: ([[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],          # short batch on rank 1
             [[3, 4, 5], [9, 10, 11], [15, 16, 17], [21, 0, 1]]),
        20: ([[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 0]],           # short batch on rank 0
             [[3, 4, 5], [9, 10, 11], [15, 16, 17], [1, 2, 3]]),
        2: ([[0, 1, 0]], [[1, 0, 1]]),                                    # smaller than one batch
    }
    for n, want in cases.items():
        base = BatchSampler(range(n), batch_size=3, drop_last=False)
        for cls in (ShardedBatchSampler, BatchSamplerShard):
            got = tuple(list(cls(base, 2, r)) for r in range(2))
            assert got == want, (n, cls.__name__, got)
    # drop_last: incomplete rounds and short batches are dropped
    base = BatchSampler(range(22), batch_size=3, drop_last=True)         # 7 full batches
    for cls in (ShardedBatchSampler, BatchSamplerShard):
        got = [list(cls(base, 2, r)) for r in range(2)]
        assert got == [[[0, 1, 2], [6, 7, 8], [12, 13, 14]], [[3, 4, 5], [9, 10, 11], [15, 16, 17]]]
