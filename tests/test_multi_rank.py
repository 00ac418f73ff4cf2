"""N > 1 host logic on CPU: world_size 2 over gloo (the GPU path uses the same functions over NCCL)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"))


def test_channel_assignment_is_a_balanced_partition():
    from lora_phy_b200.shard import assign_channels, packet_range
    rng = np.random.default_rng(0)
    sfs = rng.integers(7, 13, 4096)
    samples = [int(40 * (1 << sf)) for sf in sfs]           # configs[4]: 4096 channels, SF uniform over 7..12
    for world in (1, 2, 4, 8):
        owned = assign_channels(samples, world)
        flat = sorted(c for r in owned for c in r)
        assert flat == list(range(4096))                     # every channel exactly once
        load = [sum(samples[c] for c in r) for r in owned]
        assert max(load) - min(load) <= max(samples)         # spread bounded by one largest channel
    for n, world in ((65536, 8), (10, 4), (3, 8)):
        spans = [packet_range(n, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from lora_phy_b200.shard import gather_records, gather_records_to_root, packet_range
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 64
    lo, hi = packet_range(world * n, world, rank)
    # a rank's decoded records: [n_bytes, crc_ok, global packet number ...] in the 16-byte layout of the C ABI
    rec = torch.zeros((n, 16), dtype=torch.uint8)
    rec[:, 0] = 16
    rec[:, 11] = 1
    ids = torch.arange(lo, hi, dtype=torch.int32)
    rec[:, 12:16] = ids.view(torch.uint8).reshape(n, 4)
    out = gather_records(rec, dist)
    got = out[:, 12:16].contiguous().view(torch.int32).reshape(-1)
    ok = bool(torch.equal(got, torch.arange(world * n, dtype=torch.int32))) and int(out[:, 11].sum()) == world * n
    # the gateway's form: only the forwarding rank receives
    at_root = gather_records_to_root(rec, dist, root=0)
    ok = ok and ((at_root is None) if rank != 0 else bool(torch.equal(at_root, out)))
    # max-over-ranks timing reduction as bench.py does it
    ms = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    q.put((rank, ok, float(ms.item())))
    dist.destroy_process_group()


def test_record_gather_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[1] for r in res] == [True, True] and all(r[2] == 11.0 for r in res)


def _ragged_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from lora_phy_b200.shard import assign_channels, gather_ragged_records, ragged_counts, valid_rows
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # a small gateway: channels of very different packet counts, balanced by samples (scripts/gateway.py does the same)
    rng = np.random.default_rng(2)
    sfs = rng.integers(7, 13, 50)
    packets = {7: 91, 8: 46, 9: 27, 10: 13, 11: 8, 12: 4}
    samples = [packets[int(s)] * 40 * (1 << int(s)) for s in sfs]
    mine = assign_channels(samples, world)[rank]
    ids = [(c, k) for c in mine for k in range(packets[int(sfs[c])])]
    counts = ragged_counts(len(ids), dist)
    n_max = max(counts)
    rec = torch.zeros((n_max, 8), dtype=torch.uint8)
    rec[:len(ids)] = torch.from_numpy(np.asarray(ids, np.uint32).view(np.uint8).reshape(len(ids), 8))
    out = gather_ragged_records(rec, dist).numpy()
    seen = set()
    for r, first, n in valid_rows(counts, n_max):
        got = np.ascontiguousarray(out[first:first + n]).view(np.uint32).reshape(-1, 2)
        seen.update(map(tuple, got.tolist()))
    want = {(c, k) for c in range(50) for k in range(packets[int(sfs[c])])}
    q.put((rank, seen == want and counts[rank] == len(ids) and len(set(counts)) > 1))
    dist.destroy_process_group()


def test_ragged_record_gather_world_size_2_gloo():
    """Ranks of a gateway hold different packet counts: padded all-gather + per-rank counts must cover every
    (channel, packet) exactly once."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_ragged_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[1] for r in res] == [True, True]
