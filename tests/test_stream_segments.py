"""Host logic that lets one long capture be scanned as many independent streams: segmentation with a
frame-length overlap and the ownership rule that removes duplicates.  The synchroniser here is the CPU oracle, so the
test runs without a GPU; tests/test_gpu_stream.py shows the kernel equals the oracle stream by stream."""
import numpy as np
import pytest

from _oracle import ORC
from lora_phy_b200.shard import merge_stream_packets, stream_segments


def test_segments_partition_and_overlap():
    segs = stream_segments(1_000_003, 7, 256, mtu=40)
    assert segs[0]["own_lo"] == 0 and segs[-1]["own_hi"] == 1_000_003
    for a, b in zip(segs, segs[1:]):
        assert a["own_hi"] == b["own_lo"]                       # owned ranges tile the capture
    for s in segs:
        assert 0 <= s["lo"] <= s["own_lo"] < s["own_hi"] <= s["hi"] <= 1_000_003
        assert s["own_lo"] - s["lo"] in (0, 16 * 256) or s["lo"] == 0
        assert s["hi"] - s["own_hi"] == 42 * 256 or s["hi"] == 1_000_003
    with pytest.raises(ValueError):
        stream_segments(10, 0, 128)


@pytest.mark.parametrize("sf,n_segments", [(7, 5), (8, 3), (9, 4)])
def test_segmented_scan_finds_every_frame_once(sf, n_segments):
    n = 1 << sf
    rng = np.random.default_rng(sf)
    mtu = 24
    frames, parts = [], []
    at = 0
    for _ in range(14):
        gap = int(rng.integers(2 * n, 9 * n))
        sym = rng.integers(0, n, mtu, dtype=np.uint16)
        fr = ORC.frame_modulate(sf, sym, padding=0, ampl=1.0)
        parts += [np.zeros(gap, np.complex64), fr]
        frames.append((at + gap, sym))
        at += gap + len(fr)
    parts.append(np.zeros(4 * n, np.complex64))
    capture = np.concatenate(parts)
    capture = (capture + (rng.standard_normal(len(capture)) + 1j * rng.standard_normal(len(capture))) * 0.05).astype(np.complex64)

    segs = stream_segments(len(capture), n_segments, n, mtu=mtu)
    per_segment = []
    for s in segs:
        pk, start, err, _ = ORC.stream_demod(sf, capture[s["lo"]:s["hi"]], mtu=mtu, thresh=-10.0)
        per_segment.append([(int(start[k]), pk[k], int(err[k])) for k in range(len(pk))])
    merged = merge_stream_packets(segs, per_segment)

    # every frame exactly once, at its place, with its symbols
    assert len(merged) == len(frames)
    for (at, sym, _), (f_at, f_sym) in zip(merged, frames):
        assert abs(at - (f_at + 14 * n + n // 4)) <= 2
        np.testing.assert_array_equal(sym[:mtu].astype(np.int64), f_sym.astype(np.int64))
    # and the same packets a single sequential scan reports
    pk, start, err, _ = ORC.stream_demod(sf, capture, mtu=mtu, thresh=-10.0)
    assert [int(s) for s in start] == [m[0] for m in merged]
