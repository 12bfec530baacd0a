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


class JobPlan(object):
    """How a captioning job over `event_counts` (events per video, the whole job, known to every rank) is laid
    out on `world` ranks: rank r takes the contiguous block of videos `bounds[r]` (balanced by events) and feeds
    it in `n_batches` batches of at most `batch_videos` videos.  Every rank uses the SAME number of batches (a
    rank with fewer videos has smaller, possibly empty, ones), because each batch ends in one collective: the
    all_gather of that batch's token ids.  `batches[r][b]` = (first video, end video); `events[r][b]` = decode
    rows of that batch, i.e. the row counts `gather_tokens` needs without a count exchange.  Everything here is
    a pure function of (event_counts, world, batch_videos): no communication."""

    def __init__(self, event_counts, world, batch_videos):
        self.world = max(1, int(world))
        self.bounds = [shard_videos(event_counts, r, self.world) for r in range(self.world)]
        self.n_batches = max(1, max(-(-(hi - lo) // int(batch_videos)) for lo, hi in self.bounds))
        self.batches, self.events = [], []
        for lo, hi in self.bounds:
            cuts = [lo + ((hi - lo) * b + self.n_batches // 2) // self.n_batches for b in range(self.n_batches + 1)]
            cuts[0], cuts[-1] = lo, hi
            self.batches.append([(cuts[b], cuts[b + 1]) for b in range(self.n_batches)])
            self.events.append([int(sum(event_counts[a:b])) for a, b in self.batches[-1]])

    def counts(self, b):
        """Rows every rank contributes to the gather of batch slot b."""
        return [self.events[r][b] for r in range(self.world)]

    def assemble(self, gathered):
        """gathered[b] = the token matrix gather_tokens returned for batch slot b (rows in rank order) -> the
        job's [n_events, S] matrix in video order."""
        parts = []
        for r in range(self.world):
            for b in range(self.n_batches):
                row0 = sum(self.events[q][b] for q in range(r))
                parts.append(gathered[b][row0:row0 + self.events[r][b]])
        return torch.cat(parts, 0)


def exchange_counts(n_local, device, group=None):
    """Row count of every rank (one small all_gather + host read; do it once per batch shape)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [int(n_local)]
    world = dist.get_world_size(group)
    n = torch.tensor([int(n_local)], dtype=torch.int64, device=device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    return [int(c.item()) for c in counts]


def gather_tokens(tokens, group=None, counts=None):
    """tokens [n_local, S] (any integer/float dtype) -> [sum n_local, S] on every rank, in rank
    order.  Row counts may differ per rank: pass `counts` (from exchange_counts) to skip the count
    exchange and its host sync; the payload is padded to the largest shard and moved by ONE
    all_gather_into_tensor."""
    if not dist.is_available() or not dist.is_initialized():
        return tokens
    world = dist.get_world_size(group)
    if world == 1:
        return tokens
    if counts is None:
        counts = exchange_counts(tokens.shape[0], tokens.device, group)
    mx = max(counts)
    pad = tokens
    if tokens.shape[0] < mx:
        pad = torch.cat([tokens, tokens.new_zeros((mx - tokens.shape[0],) + tuple(tokens.shape[1:]))], 0)
    out = torch.empty((world * mx,) + tuple(tokens.shape[1:]), dtype=tokens.dtype, device=tokens.device)
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    if all(c == mx for c in counts):
        return out
    return torch.cat([out[r * mx:r * mx + c] for r, c in enumerate(counts)], 0)


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
