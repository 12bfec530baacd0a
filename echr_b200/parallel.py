"""Multi-GPU plumbing: videos are independent, so a batch is sharded by video with
no data-path collective; the only exchange is one all_gather of the decoded token ids (and,
optionally, their log-probs) at the end of a batch.  One process per GPU, torch.distributed
(NCCL over NVLink on the GPU box; gloo in the CPU tests)."""
import torch
import torch.distributed as dist


def shard_videos(event_counts, rank, world):
    """Contiguous block of videos for `rank`, balanced by number of events (decode cost is
    proportional to events).  Returns (first_video, end_video)."""
    n = len(event_counts)
    if world <= 1:
        return 0, n
    total = float(sum(event_counts)) or 1.0
    bounds, acc, nxt = [0], 0.0, 1
    for i, c in enumerate(event_counts):
        acc += c
        while nxt < world and acc >= total * nxt / world:
            bounds.append(i + 1)
            nxt += 1
    while len(bounds) < world:
        bounds.append(n)
    bounds.append(n)
    # keep bounds monotone and leave no rank without a video when there are enough videos
    for r in range(1, world + 1):
        bounds[r] = max(bounds[r], bounds[r - 1])
    return bounds[rank], bounds[rank + 1]


def gather_tokens(tokens, group=None):
    """tokens [n_local, S] (any integer/float dtype) -> [sum n_local, S] on every rank, in rank
    order.  Row counts may differ per rank: they are exchanged first and the payload is padded to
    the largest shard (one small + one payload collective)."""
    if not dist.is_available() or not dist.is_initialized():
        return tokens
    world = dist.get_world_size(group)
    if world == 1:
        return tokens
    n = torch.tensor([tokens.shape[0]], dtype=torch.int64, device=tokens.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    mx = max(counts)
    pad = tokens
    if tokens.shape[0] < mx:
        pad = torch.cat([tokens, tokens.new_zeros((mx - tokens.shape[0],) + tuple(tokens.shape[1:]))], 0)
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad.contiguous(), group=group)
    return torch.cat([o[:c] for o, c in zip(out, counts)], 0)


def caption_sharded(model, videos, beam_size=1, group=None):
    """Caption `videos` (the same list on every rank) with each rank decoding its shard; every rank
    returns the full [n_events, S] token matrix, identical to a single-GPU run."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    counts = [len(v[4]) for v in videos]
    lo, hi = shard_videos(counts, rank, world)
    eng = model.engine
    S = eng.seq_length
    if hi > lo:
        eng_, layout = model._stage(videos[lo:hi], max(1, beam_size))
        tok = eng.decode_beam(beam_size)[0] if beam_size > 1 else eng.decode_greedy()[0]
    else:
        tok = torch.zeros(0, S, dtype=torch.int32, device=eng.device)
    return gather_tokens(tok, group)
