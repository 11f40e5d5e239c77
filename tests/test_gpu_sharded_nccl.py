"""GPU, 2 ranks over NCCL (skipped on a one-GPU box): generate_sharded() - every rank generates its contiguous batch
shard on its own GPU with NO per-step collective, the audio is all-gathered at the end - equals the unsharded
generate() of the global batch sample for sample, with top-k SAMPLING: the Philox counter is keyed by the global row
(row_offset), so the result does not depend on the number of GPUs (the caller pattern of
/root/reference/training/run_parler_tts_training.py:1548-1556)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import parler_tts_b200 as ptts
        from parler_tts_b200 import sharding
        from test_gpu_generate import _composite

        model, dims, sd, dac_sd = _composite()
        model = model.to(f"cuda:{rank}")
        g = torch.Generator().manual_seed(5)
        B = 5  # shards of 3 and 2
        input_ids, prompt_ids = torch.randint(2, 100, (B, 7), generator=g), torch.randint(2, 100, (B, 4), generator=g)
        kw = dict(do_sample=True, top_k=12, seed=21, max_length=22, logits_processor=[ptts.SuppressTokensAtOrAbove(32)])
        dev = f"cuda:{rank}"
        whole = model.generate(input_ids=input_ids.to(dev), prompt_input_ids=prompt_ids.to(dev), **kw)
        sharded = sharding.generate_sharded(model, input_ids.to(dev), prompt_ids.to(dev), **kw)
        np.save(os.path.join(out_dir, f"whole{rank}.npy"), whole.cpu().numpy())
        np.save(os.path.join(out_dir, f"sharded{rank}.npy"), sharded.cpu().numpy())
    finally:
        dist.destroy_process_group()


def test_world2_nccl_sharded_sampling_equals_unsharded(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    w0, s0, s1 = (np.load(tmp_path / f) for f in ("whole0.npy", "sharded0.npy", "sharded1.npy"))
    assert s0.shape == w0.shape and s0.shape[0] == 5
    assert np.array_equal(s0, s1)            # every rank holds the gathered result
    np.testing.assert_allclose(s0, w0, atol=1e-6)  # same tokens -> same audio, whichever GPU produced the row
    assert np.abs(w0).max() > 0
