"""CPU, world_size 2 over gloo: the multi-GPU path's host logic (contiguous batch shards, global row offsets, padded
gather of finished codes / audio).  The per-step path has no collective."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from parler_tts_b200 import sharding


def test_shard_bounds_cover_the_batch_contiguously():
    for total in (0, 1, 7, 64, 512, 513):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        ids = torch.randint(0, 100, (5, 7), generator=g)         # global batch of 5 utterances: shards of 3 and 2
        prm = torch.randint(0, 100, (5, 3), generator=g)
        (a, b, c), start = sharding.shard_batch((ids, prm, None))
        lo, hi = sharding.shard_bounds(5, world, rank)
        assert start == lo and torch.equal(a, ids[lo:hi]) and torch.equal(b, prm[lo:hi]) and c is None
        # ragged "audio": rank r produced utterances of length 4 + 3 r; codes (B_r, K, F_r)
        audio = torch.arange((hi - lo) * (4 + 3 * rank), dtype=torch.float32).reshape(hi - lo, -1) + 100 * rank + 1
        codes = torch.full((hi - lo, 9, 2 + rank), rank + 1, dtype=torch.int64)
        ga = sharding.gather_padded(audio, 0.0)
        gc = sharding.gather_padded(codes, 1024)
        np.save(os.path.join(out_dir, f"audio{rank}.npy"), ga.numpy())
        np.save(os.path.join(out_dir, f"codes{rank}.npy"), gc.numpy())
    finally:
        dist.destroy_process_group()


def test_world2_gloo_shard_and_padded_gather(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    a0, a1 = np.load(tmp_path / "audio0.npy"), np.load(tmp_path / "audio1.npy")
    c0, c1 = np.load(tmp_path / "codes0.npy"), np.load(tmp_path / "codes1.npy")
    assert np.array_equal(a0, a1) and np.array_equal(c0, c1)      # every rank holds the full result
    assert a0.shape == (5, 7) and c0.shape == (5, 9, 3)
    assert np.array_equal(a0[:3, :4], np.arange(12, dtype=np.float32).reshape(3, 4) + 1)
    assert np.all(a0[:3, 4:] == 0)                                 # zero right-padding like pad_sequence(..., 0)
    assert np.array_equal(a0[3:], np.arange(14, dtype=np.float32).reshape(2, 7) + 101)
    assert np.all(c0[:3, :, :2] == 1) and np.all(c0[:3, :, 2] == 1024) and np.all(c0[3:] == 2)


class _FakeModel:
    """Stands in for the composite model on CPU: `generate` returns one audio row per utterance whose content encodes the
    description ids and the global row offset it was called with."""
    device = torch.device("cpu")

    def generate(self, input_ids, prompt_input_ids, row_offset=0, **kw):
        B = input_ids.shape[0]
        assert B > 0, "an empty shard must not reach generate()"
        T = 3 + int(row_offset)  # ragged across ranks
        rows = (input_ids[:, :1].float() * 10 + row_offset + torch.arange(B)[:, None]).repeat(1, T)
        if "attention_mask" in kw:
            rows = rows + 0.5 * kw["attention_mask"].sum(dim=1, keepdim=True)
        return rows


def _worker_generate(rank, world, port, out_dir, n_utt):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ids = torch.arange(1, n_utt + 1)[:, None].repeat(1, 4)
        prm = torch.zeros(n_utt, 2, dtype=torch.int64)
        am = torch.ones(n_utt, 4, dtype=torch.int64)
        out = sharding.generate_sharded(_FakeModel(), ids, prm, attention_mask=am)
        np.save(os.path.join(out_dir, f"gen{rank}.npy"), out.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_utt", [1, 3])
def test_world2_gloo_generate_sharded_incl_an_empty_shard(tmp_path, n_utt):
    """The eval-side caller pattern (training/run_parler_tts_training.py:1548-1556) over two ranks: every rank returns the
    whole batch, zero padded to the longest shard's audio; with one utterance rank 1 holds an EMPTY shard, skips
    generate() and still takes part in the gather (no hang, no error)."""
    world, port = 2, _free_port()
    mp.spawn(_worker_generate, args=(world, port, str(tmp_path), n_utt), nprocs=world, join=True)
    g0, g1 = np.load(tmp_path / "gen0.npy"), np.load(tmp_path / "gen1.npy")
    assert np.array_equal(g0, g1) and g0.shape[0] == n_utt
    if n_utt == 1:
        assert g0.shape == (1, 3) and np.all(g0 == 10 + 0 + 0 + 2.0)
    else:  # shards of 2 (offset 0, T = 3) and 1 (offset 2, T = 5)
        assert g0.shape == (3, 5)
        assert np.all(g0[0, :3] == 10 + 2.0) and np.all(g0[1, :3] == 20 + 1 + 2.0) and np.all(g0[:2, 3:] == 0)
        assert np.all(g0[2] == 30 + 2 + 2.0)
