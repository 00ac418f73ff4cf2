"""Multi-GPU plumbing of the receive path: which rank owns which channel, and the one collective of a step.

Packets / channels / streams are independent: every rank demodulates and decodes its own
share with no data-path collective, and only the fixed-stride decoded records cross NVLink -- one gather to the
forwarding rank per step (or an all-gather where every rank wants them; NCCL on GPUs, the same code runs on gloo with
CPU tensors in the tests).
"""


def assign_channels(samples_per_channel, world):
    """Static balanced partition for BASELINE configs[4]: channels have very different costs (an SF12 channel
    carries 32x the samples of an SF7 one), so sort by sample count and give each channel to the least-loaded
    rank (longest-processing-time rule).  -> list of channel-index lists, one per rank; every channel appears
    exactly once; the load spread is at most one largest channel."""
    order = sorted(range(len(samples_per_channel)), key=lambda c: (-samples_per_channel[c], c))
    load = [0] * world
    owned = [[] for _ in range(world)]
    for c in order:
        r = min(range(world), key=lambda k: (load[k], k))
        owned[r].append(c)
        load[r] += samples_per_channel[c]
    return owned


def packet_range(n_packets, world, rank):
    """Contiguous split of a packet batch (configs[1]/[2]): rank r owns [lo, hi)."""
    lo = n_packets * rank // world
    hi = n_packets * (rank + 1) // world
    return lo, hi


def gather_records(local, dist, out=None):
    """All-gather of per-packet records.  `local`: [n, stride] uint8 tensor (same n on every rank), on the device the
    process group's backend works with.  -> [world * n, stride], rank-major, identical on every rank."""
    import torch
    world = dist.get_world_size()
    if out is None:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous())
    return out


def gather_records_to_root(local, dist, root=0, out=None):
    """The gateway's collective: every rank's records travel ONCE, to the rank that forwards them (`root`), instead of
    being replicated on all ranks -- 1/world of the all-gather's bytes over NVLink.  `local`: [n, stride] uint8, same n
    on every rank.  -> [world * n, stride] rank-major on `root` (written into `out` when given), None elsewhere."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    local = local.contiguous()
    if rank != root:
        dist.gather(local, gather_list=None, dst=root)
        return None
    if out is None:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    n = local.shape[0]
    dist.gather(local, gather_list=[out[r * n:(r + 1) * n] for r in range(world)], dst=root)
    return out


def ragged_counts(n_local, dist, device=None):
    """Record counts of every rank (a gateway's ranks own different numbers of packets: channels are balanced by
    samples, not by packets).  One small all-gather at set-up time.  -> list of ints, rank order."""
    import torch
    mine = torch.tensor([int(n_local)], dtype=torch.int64, device=device)
    parts = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return [int(p.item()) for p in parts]


def gather_ragged_records(local_padded, dist, out=None, root=None):
    """Gather of per-packet records when ranks hold different counts: every rank passes a buffer of the SAME padded
    length (max over `ragged_counts`), rows past its own count are ignored by the reader.  root=None: all-gather
    (identical on every rank); root=r: only rank r receives (gather_records_to_root).  -> [world * n_max, stride];
    use `valid_rows` to walk it."""
    if root is not None:
        return gather_records_to_root(local_padded, dist, root=root, out=out)
    return gather_records(local_padded, dist, out=out)


def valid_rows(counts, n_max):
    """(rank, first_row, n_rows) of every rank's valid records inside a gather_ragged_records result."""
    return [(r, r * n_max, c) for r, c in enumerate(counts)]


def stream_segments(count, n_segments, n, mtu=256, preamble_symbols=16):
    """Cut one long capture of `count` samples into `n_segments` pieces for the frame synchroniser (SURVEY.md 8e,
    "overlap-save by one maximum frame").  Segment k OWNS [own_lo, own_hi) -- a partition of [0, count) -- and READS
    [lo, hi): `preamble_symbols*n` samples earlier, so a frame whose data starts just inside the owned range still has
    its whole preamble (10 up-chirps, sync word, 2 1/4 down-chirps) in view, and `(mtu + 2)*n` samples later, so a
    packet that starts at the end of the owned range is seen to its last symbol.  -> list of dicts."""
    if n_segments < 1 or count < 0:
        raise ValueError("bad segmentation")
    out = []
    for k in range(n_segments):
        own_lo = count * k // n_segments
        own_hi = count * (k + 1) // n_segments
        lo = max(0, own_lo - preamble_symbols * n)
        hi = min(count, own_hi + (mtu + 2) * n)
        out.append({"own_lo": own_lo, "own_hi": own_hi, "lo": lo, "hi": hi})
    return out


def merge_stream_packets(segments, per_segment):
    """per_segment[k] = [(start_in_segment, symbols, freq_error), ...] as the synchroniser reported them for the
    samples [lo, hi) of segment k.  A packet is kept from the segment that owns its first data sample, which removes
    the copies neighbouring segments see in their overlap.  -> [(absolute_start, symbols, freq_error)] in stream order."""
    merged = []
    for seg, packets in zip(segments, per_segment):
        for start, symbols, err in packets:
            at = seg["lo"] + int(start)
            if seg["own_lo"] <= at < seg["own_hi"]:
                merged.append((at, symbols, err))
    merged.sort(key=lambda p: p[0])
    return merged
