"""Multi-GPU plumbing of the receive path: which rank owns which channel, and the one collective of a step.

Packets / channels / streams are independent: every rank demodulates and decodes its own
share with no data-path collective, and only the fixed-stride decoded records cross NVLink -- one
all_gather_into_tensor per step (NCCL on GPUs; the same code runs on gloo with CPU tensors in the tests).
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
