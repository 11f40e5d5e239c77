"""Batch sharding of independent utterances across the GPUs of one box.

Utterances never interact (attention, KV cache, EOS flags and DAC decode are per sequence), so the generation path
shards with NO collective on the per-step path: every rank owns a contiguous slice of the batch, replays its own CUDA
graph and stops on its own flag.  The only exchange is the gather of the finished codes / audio, padded on the last
dimension to the longest rank like the reference trainer's `pad_across_processes(dim=1, pad_index=0)` +
`gather_for_metrics` (/root/reference/training/run_parler_tts_training.py:1555) and `pad_sequence(padding_value=0)`
(/root/reference/parler_tts/modeling_parler_tts.py:2676-2677).  Works on NCCL (device tensors) and gloo (CPU tensors).
"""
import torch
import torch.distributed as dist


def shard_bounds(total, world, rank):
    """Contiguous split of `total` items over `world` ranks (first `total % world` ranks get one more)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} of {world}")
    base, extra = divmod(total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_batch(tensors, world=None, rank=None):
    """Slices every tensor of `tensors` (dict or tuple, batch on dim 0) to this rank's utterances.
    Returns (sharded, global_row_offset): pass the offset to generate(row_offset=...) so that sampled tokens do not
    depend on the number of GPUs (the Philox counter is keyed by the global row index)."""
    world = dist.get_world_size() if world is None else world
    rank = dist.get_rank() if rank is None else rank
    items = tensors.items() if isinstance(tensors, dict) else enumerate(tensors)
    out, start = {}, 0
    for k, t in items:
        if t is None:
            out[k] = None
            continue
        start, stop = shard_bounds(t.shape[0], world, rank)
        out[k] = t[start:stop]
    if not isinstance(tensors, dict):
        out = tuple(out[i] for i in range(len(out)))
    return out, start


def gather_padded(x, pad_value=0, group=None):
    """All-gathers per-rank results of shape (B_r, ..., T_r): pads the last dim to the longest rank with `pad_value`,
    the batch dim to the largest shard, gathers, and returns the concatenation (sum B_r, ..., max T_r) on every rank."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return x
    world = dist.get_world_size(group)
    shape = torch.tensor([x.shape[0], x.shape[-1]], dtype=torch.int64, device=x.device)
    shapes = [torch.empty_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    shapes = [tuple(int(v) for v in s.tolist()) for s in shapes]
    bmax, tmax = max(max(s[0] for s in shapes), 1), max(s[1] for s in shapes)
    buf = torch.full((bmax,) + tuple(x.shape[1:-1]) + (tmax,), pad_value, dtype=x.dtype, device=x.device)
    buf[: x.shape[0], ..., : x.shape[-1]] = x
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf.contiguous(), group=group)
    return torch.cat([p[: s[0]] for p, s in zip(parts, shapes)], dim=0)


@torch.no_grad()
def generate_sharded(model, input_ids, prompt_input_ids, attention_mask=None, prompt_attention_mask=None, **kwargs):
    """generate() over a global batch: this rank generates its contiguous shard on its own GPU, then the audio is
    gathered (zero padded to the longest utterance) so that every rank returns the full (B, T_max) result."""
    (ids, prm, am, pm), start = shard_batch((input_ids, prompt_input_ids, attention_mask, prompt_attention_mask))
    if am is not None:
        kwargs["attention_mask"] = am
    if pm is not None:
        kwargs["prompt_attention_mask"] = pm
    if ids.shape[0] == 0:
        # more ranks than utterances: this rank has nothing to generate but must still take part in the gather
        audio = torch.zeros(0, 1, dtype=torch.float32, device=model.device)
    else:
        audio = model.generate(input_ids=ids, prompt_input_ids=prm, row_offset=start, **kwargs)
    return gather_padded(audio, 0.0)
