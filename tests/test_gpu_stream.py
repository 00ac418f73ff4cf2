"""GPU parity of the continuous-stream front end: the CUDA frame generator and frame
synchroniser, called through the C ABI, against oracle/stream.c on the same inputs.

Bar: generated IQ bit-exact (the kernel evaluates cosf/sinf with glibc's bits); symbols, packet boundaries, frequency
errors and consumed-sample counts bit-exact -- they are integer outputs of a state machine whose every decision the
kernel reproduces with the reference's float operations, including the int->float->int fine-tune index walk.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import lora_phy_b200 as L
    from _oracle import ORC
    return {"torch": torch, "L": L, "ORC": ORC}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def channel(frame, n, lead, tail, cfo_bins, snr_db, rng, ampl=0.3):
    x = np.concatenate([np.zeros(lead, np.complex64), frame, np.zeros(tail, np.complex64)])
    k = np.arange(len(x))
    x = x * np.exp(2j * np.pi * cfo_bins * k / n)
    sigma = ampl * 10 ** (-snr_db / 20) / np.sqrt(2)
    x = x + (rng.standard_normal(len(x)) + 1j * rng.standard_normal(len(x))) * sigma
    return x.astype(np.complex64)


def compare(env, sf, streams, **kw):
    """Runs the same streams (rows, equal length) through the kernel and the oracle; asserts every output equal."""
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    ctx = L.Context(sf)
    x = np.stack(streams)
    mtu = kw.get("mtu", 256)
    cap_p, cap_s = kw.get("max_packets", 8), kw.get("max_symbols", 2048)
    sym, pk, res = ctx.stream_detect_batch(torch.from_numpy(x).cuda(), sync=kw.get("sync", 0x12),
                                           thresh_db=kw.get("thresh", -30.0), mtu=mtu, max_packets=cap_p, max_symbols=cap_s)
    torch.cuda.synchronize()
    sym = sym.cpu().numpy()
    pk, res = L.stream_records(pk, res)
    found = 0
    for i, s in enumerate(streams):
        want, start, err, consumed = ORC.stream_demod(sf, s, sync=kw.get("sync", 0x12), thresh=kw.get("thresh", -30.0),
                                                      mtu=mtu, pkt_cap=cap_p)
        assert int(res[i]["consumed"]) == consumed, (i, res[i], consumed)
        assert int(res[i]["n_packets"]) == len(want), (i, res[i], len(want))
        off = 0
        for k, w in enumerate(want):
            assert int(pk[i, k]["first_symbol"]) == off and int(pk[i, k]["n_symbols"]) == len(w)
            assert int(pk[i, k]["start"]) == int(start[k]) and int(pk[i, k]["freq_error"]) == int(err[k])
            assert bits(sym[i, off:off + len(w)]) == bits(w), (i, k)
            off += len(w)
        assert int(res[i]["n_symbols"]) == off
        found += len(want)
    return found


@pytest.mark.parametrize("sf,ovs", [(7, 1), (8, 2), (9, 1), (10, 1), (12, 1)])
def test_frame_generator_bit_exact(env, sf, ovs):
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    rng = np.random.default_rng(sf)
    n = 1 << sf
    sym = rng.integers(0, n, (37, 6), dtype=np.uint16)  # 37 frames: a ragged last warp
    ctx = L.Context(sf, osr=ovs)
    iq = ctx.frame_modulate_batch(torch.from_numpy(sym.astype(np.int16)).cuda(), amplitude=0.3, sync=0x34, padding=2)
    torch.cuda.synchronize()
    iq = iq.cpu().numpy()
    assert iq.shape[1] == ctx.frame_samples(6, 2) == (14 + 6 + 2) * n * ovs + n * ovs // 4
    for p in (0, 17, 36):
        assert bits(iq[p]) == bits(ORC.frame_modulate(sf, sym[p], sync=0x34, padding=2, ampl=0.3, ovs=ovs)), p


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_stream_detect_matches_oracle(env, sf):
    ORC = env["ORC"]
    n = 1 << sf
    rng = np.random.default_rng(100 + sf)
    nsym = 12
    streams = []
    cases = [(0.0, 0, 80.0), (0.3, 37, 20.0), (-2.4, n // 2 + 3, 10.0), (5.2, 11, 0.0), (0.49, n - 1, 15.0), (-0.51, 1, 6.0)]
    for cfo, lead, snr in cases:
        sym = rng.integers(0, n, nsym, dtype=np.uint16)
        fr = ORC.frame_modulate(sf, sym, padding=3)
        x = channel(fr, n, 3 * n + lead, 0, cfo, snr, rng)
        total = (3 + 14 + nsym + 3 + 3) * n + n // 4 + n  # common length
        streams.append(np.concatenate([x, np.zeros(total - len(x), np.complex64)])
                       + ((rng.standard_normal(total) + 1j * rng.standard_normal(total)) * 1e-4).astype(np.complex64))
    assert compare(env, sf, streams, mtu=nsym) >= len(cases) - 1


def test_stream_detect_many_streams_two_frames_each(env):
    """More streams than slots in one wave, two frames per stream ended by squelch, different content per stream."""
    ORC = env["ORC"]
    sf, n = 7, 128
    rng = np.random.default_rng(3)
    streams = []
    for i in range(300):
        a = rng.integers(0, n, 9, dtype=np.uint16)
        b = rng.integers(0, n, 5, dtype=np.uint16)
        x = np.concatenate([channel(ORC.frame_modulate(sf, a, padding=3), n, 2 * n + int(rng.integers(0, n)), 0,
                                    float(rng.uniform(-1, 1)), 25.0, rng),
                            channel(ORC.frame_modulate(sf, b, padding=3), n, int(rng.integers(0, n)), 0,
                                    float(rng.uniform(-1, 1)), 25.0, rng)])
        streams.append(x)
    size = max(len(s) for s in streams) + 2 * n
    streams = [np.concatenate([s, ((rng.standard_normal(size - len(s)) + 1j * rng.standard_normal(size - len(s))) * 1e-3)
                               .astype(np.complex64)]) for s in streams]
    assert compare(env, sf, streams, mtu=64, thresh=-10.0) >= 550


@pytest.mark.parametrize("sf", [7, 9, 10])
def test_stream_detect_never_squelched_walk(env, sf):
    """thresh = -200 dB: no window is squelched, so the fine-tune error accumulates an arbitrary fractional value every
    window of noise and the int->float->int index walk runs through binade crossings, wraps and negative strides."""
    n = 1 << sf
    rng = np.random.default_rng(sf)
    streams = [((rng.standard_normal(60 * n) + 1j * rng.standard_normal(60 * n)) * 0.2).astype(np.complex64) for _ in range(12)]
    compare(env, sf, streams, thresh=-200.0, mtu=16, max_packets=16)


def test_stream_detect_edge_cases(env):
    torch, L = env["torch"], env["L"]
    sf, n = 8, 256
    rng = np.random.default_rng(0)
    noise = ((rng.standard_normal(10 * n) + 1j * rng.standard_normal(10 * n)) * 0.1).astype(np.complex64)
    compare(env, sf, [noise[: 2 * n - 1]])          # shorter than 2N: nothing consumed (LoRaDemod.cpp:148)
    compare(env, sf, [noise[: 2 * n]])              # exactly one call
    compare(env, sf, [noise], thresh=50.0)          # everything squelched
    compare(env, sf, [np.zeros(6 * n, np.complex64)])  # all-zero input: log10(0) = -inf on both sides
    # capacity overflow: more packets / symbols than the caller made room for are counted but not stored
    from _oracle import ORC
    a = rng.integers(0, n, 10, dtype=np.uint16)
    x = np.concatenate([channel(ORC.frame_modulate(sf, a, padding=3), n, 2 * n, 0, 0.0, 30.0, rng) for _ in range(3)]
                       + [np.zeros(2 * n, np.complex64)])
    ctx = L.Context(sf)
    sym, pk, res = ctx.stream_detect_batch(torch.from_numpy(x[None]).cuda(), mtu=10, max_packets=2, max_symbols=15)
    pk, res = L.stream_records(pk, res)
    assert int(res[0]["n_packets"]) == 3 and int(res[0]["n_symbols"]) == 15
    assert bits(sym.cpu().numpy()[0, :10]) == bits(a.astype(np.int16))
    # an oversampled stream is decimated on the fly: a context with osr 2 sees samples 0, 2, 4, ... and reports
    # positions in those units, exactly as the oracle does on x[::2]
    y = np.repeat(x, 2)[: 2 * len(x) - 1]                 # odd length: the last decimated sample is the very last one
    sym2, pk2, res2 = L.Context(sf, osr=2).stream_detect_batch(torch.from_numpy(y[None]).cuda(), mtu=10)
    want, start, err, consumed = ORC.stream_demod(sf, y[::2], mtu=10)
    pk2, res2 = L.stream_records(pk2, res2)
    assert int(res2[0]["n_packets"]) == len(want) == 3 and int(res2[0]["consumed"]) == consumed
    assert [int(p["start"]) for p in pk2[0, :3]] == [int(v) for v in start]
    assert bits(sym2.cpu().numpy()[0, :30]) == bits(np.concatenate(want))


def test_stream_detect_repeatable_under_load(env):
    """Same purpose as test_repeatable_under_load in test_gpu_parity.py, for the stream kernel: every slot busy, many
    rounds of the index walk (which synchronises through shared memory), identical results run after run."""
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    sf, n = 10, 1024
    rng = np.random.default_rng(17)
    frames = []
    for i in range(8):
        sym = rng.integers(0, n, 14, dtype=np.uint16)
        frames.append(channel(ORC.frame_modulate(sf, sym, padding=2, ampl=1.0), n, 2 * n + int(rng.integers(0, n)), n,
                              float(rng.uniform(-200, 200)), 12.0, rng, ampl=1.0))
    size = max(len(f) for f in frames)
    base = np.stack([np.concatenate([f, np.zeros(size - len(f), np.complex64)]) for f in frames])
    x = torch.from_numpy(np.tile(base, (160, 1))).cuda()          # 1280 streams: more than one wave of slots
    ctx = L.Context(sf)
    first = None
    for rep in range(6):
        sym, pk, res = ctx.stream_detect_batch(x, mtu=14, thresh_db=-8.0)
        cur = (sym.clone(), pk.clone(), res.clone())
        if first is None:
            first = cur
        else:
            for a, b in zip(first, cur):
                assert torch.equal(a, b), rep
    sym, pk, res = first
    for k in range(8):                                             # every copy of a stream gives the same answer
        for a in (sym, pk, res):
            blk = a[k::8]
            assert (blk == blk[0]).all()
    want, start, err, consumed = ORC.stream_demod(sf, base[3], mtu=14, thresh=-8.0)
    pk_h, res_h = L.stream_records(pk, res)
    assert int(res_h[3]["n_packets"]) == len(want) and int(res_h[3]["consumed"]) == consumed


@pytest.mark.parametrize("sf,rdd,ppm", [(7, 1, 0), (9, 4, 0), (10, 2, 8)])
def test_stream_to_payload(env, sf, rdd, ppm):
    """The whole continuous-stream receiver: payload -> LoRaEncoder -> LoRaMod frame -> channel -> LoRaDemod ->
    LoRaDecoder -> payload, all on the device, two frames per stream; decoder results equal the oracle's on the
    symbols the synchroniser delivered."""
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    from _oracle import chain_cfg as orc_cfg
    n = 1 << sf
    ctx = L.Context(sf)
    cfg, oc = L.chain_cfg(sf, rdd, ppm=ppm), orc_cfg(sf, rdd, ppm=ppm)
    gen = torch.Generator(device="cuda").manual_seed(sf)
    n_streams = 40
    pay = torch.randint(0, 256, (2 * n_streams, 24), dtype=torch.uint8, device="cuda", generator=gen)
    sym = ctx.chain_encode_batch(cfg, pay)
    nsym = sym.shape[1]
    fr = ctx.frame_modulate_batch(sym, amplitude=1.0, padding=3)
    lead = 3 * n
    x = torch.zeros((n_streams, lead + 2 * (fr.shape[1] + n) + 2 * n), dtype=torch.complex64, device="cuda")
    x[:, lead:lead + fr.shape[1]] = fr[0::2]
    x[:, lead + fr.shape[1] + n: lead + 2 * fr.shape[1] + n] = fr[1::2]
    ctx.awgn_batch(x, torch.full((n_streams,), 12.0, device="cuda"), seed=sf)
    d_sym, d_pk, d_res = ctx.stream_detect_batch(x, mtu=nsym, thresh_db=-10.0, max_packets=4, max_symbols=4 * nsym)
    payload, cres = ctx.stream_decode_batch(cfg, d_sym, d_pk, d_res)
    pk, res = L.stream_records(d_pk, d_res)
    cres = cres.cpu().numpy().view(L.CHAIN_RESULT).reshape(n_streams, 4)
    payload, d_sym, pay_h = payload.cpu().numpy(), d_sym.cpu().numpy(), pay.cpu().numpy()
    good = 0
    for i in range(n_streams):
        assert int(res[i]["n_packets"]) >= 2
        for k in range(4):
            if k >= int(res[i]["n_packets"]):
                assert int(cres[i, k]["n_bytes"]) == -1
                continue
            s0, cnt = int(pk[i, k]["first_symbol"]), int(pk[i, k]["n_symbols"])
            ro, bo, info = ORC.chain_decode(oc, d_sym[i, s0:s0 + cnt].astype(np.uint16))
            assert int(cres[i, k]["n_bytes"]) == ro and int(cres[i, k]["crc_ok"]) == int(info.crc_ok)
            assert bits(payload[i, k, :max(ro, 0)]) == bits(bo)
        for k in range(2):
            if int(cres[i, k]["crc_ok"]) == 1:
                assert bytes(payload[i, k, :24]) == bytes(pay_h[2 * i + k])
                good += 1
    assert good >= 2 * n_streams - 2
