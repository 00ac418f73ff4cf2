"""lora_b200_demod_batch_static -- the receive path for callers that may not allocate -- against lora_b200_demod_batch.

The entry point has no counterpart in the reference (include/lora_phy_b200.h says who asked for it); its contract is
"the same results as lora_b200_demod_batch, bit for bit, with the scratch in a fixed device array and no allocation".
So the check is equality with the existing entry point through the C ABI, on two calls in a row that share the array
(the second one different in size and content, so anything left behind by the first would show), plus the size limits
and the no-allocation claim itself (free device memory does not move across the calls).

This file sorts last on purpose: a failure here must not hide the parity tests.
"""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import lora_phy_b200 as L
    return {"torch": torch, "L": L}


def bits(t):
    return np.ascontiguousarray(t.cpu().numpy()).view(np.uint8).tobytes()


def noisy_packets(torch, ctx, sf, packets, nsym, seed):
    n = 1 << sf
    g = torch.Generator(device="cuda").manual_seed(seed)
    sym = torch.randint(0, n, (packets, nsym), dtype=torch.int16, device="cuda", generator=g)
    iq = ctx.modulate_batch(sym)
    snr = (torch.rand(packets, device="cuda", generator=g) * 30.0 - 18.0).float().contiguous()
    ctx.awgn_batch(iq, snr, seed=seed)
    return iq.contiguous()


@pytest.mark.gpu
@pytest.mark.parametrize("sf,mode_name", [(7, "workspace"), (9, "legacy"), (12, "dechirp")])
def test_static_scratch_entry_point_equals_demod_batch_on_two_calls_in_a_row(env, sf, mode_name):
    torch, L = env["torch"], env["L"]
    mode = {"workspace": L.DEMOD_WORKSPACE, "legacy": L.DEMOD_LEGACY, "dechirp": L.DEMOD_LEGACY_DECHIRP}[mode_name]
    ctx = L.Context(sf)
    # two batches of different shape and content; the second is the larger one, so the grow-only buffers of the
    # existing entry point WOULD have to grow for it
    shapes = [(96, 20, 1), (1024 if sf < 12 else 160, 33, 2)]
    inputs = [noisy_packets(torch, ctx, sf, p, nsym, seed=100 * sf + k) for p, nsym, k in shapes]
    want = []
    for iq in inputs:
        s, m, pk = ctx.demod_batch(iq, mode=mode, want_peaks=True)
        torch.cuda.synchronize()
        want.append((bits(s), bits(m), pk.clone()))
    outs = []
    for iq in inputs:  # every buffer the calls need exists before the first of them
        p, total = iq.shape[0], iq.shape[1] // ctx.step
        outs.append((torch.zeros((p, total - 2), dtype=torch.int16, device="cuda"),
                     torch.zeros((p, 16), dtype=torch.uint8, device="cuda"),
                     torch.zeros((p, total), dtype=torch.float32, device="cuda")))
    torch.cuda.synchronize()
    free = []
    for iq, (s, m, pk) in zip(inputs, outs):
        ctx.demod_batch_static(iq, s, metrics=m, peaks=pk, mode=mode)
        torch.cuda.synchronize()
        free.append(torch.cuda.mem_get_info()[0])
    for k, ((s, m, pk), (ws, wm, wpk)) in enumerate(zip(outs, want)):
        assert bits(s) == ws, (sf, mode_name, k, "symbols")
        assert bits(m) == wm, (sf, mode_name, k, "metrics")
        # the same kernels on the same inputs: the peaks are the same floats, well inside the 1e-4 relative
        # tolerance BASELINE.json allows the fused flavour against the reference
        a, b = pk.cpu().numpy(), wpk.cpu().numpy()
        assert ((a == b) | (np.isnan(a) & np.isnan(b))).all(), (sf, mode_name, k, "peaks")
    assert free[1] == free[0], "the second, larger call changed the amount of free device memory"
    # and once more over the first batch: the array now holds the larger call's leftovers
    s, m, pk = outs[0]
    s.zero_(); m.zero_()
    ctx.demod_batch_static(inputs[0], s, metrics=m, peaks=pk, mode=mode)
    torch.cuda.synchronize()
    assert bits(s) == want[0][0] and bits(m) == want[0][1]


@pytest.mark.gpu
def test_static_scratch_entry_point_rejects_what_does_not_fit(env):
    torch, L = env["torch"], env["L"]
    ctx = L.Context(7)
    lib = L.load_library()
    buf = torch.zeros(4096, dtype=torch.complex64, device="cuda")
    sym = torch.zeros(4096, dtype=torch.int16, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def call(n_packets, count, h=ctx.h):
        # sizes are checked before anything is read: a small buffer stands in for the oversized batch
        return lib.lora_b200_demod_batch_static(h, L.DEMOD_WORKSPACE, C.c_void_p(buf.data_ptr()), n_packets, count, count,
                                                C.c_void_p(sym.data_ptr()), max(count // 128, 2), None, None, stream)

    assert call(16385, 4 * 128) != 0
    assert b"LORA_B200_STATIC_MAX_PACKETS" in lib.lora_b200_last_error()
    assert call(16384, 65 * 128) != 0                       # 16384 x 65 = 1,064,960 windows > 1,048,576
    assert b"LORA_B200_STATIC_MAX_WINDOWS" in lib.lora_b200_last_error()
    assert call(0, 4 * 128) == 0                            # an empty batch is fine, as for lora_b200_demod_batch
    small = L.Context(5)
    assert call(4, 4 * 32, h=small.h) != 0                  # SF < 7 has no blocked kernels
    assert b"sf 7..12" in lib.lora_b200_last_error()
    torch.cuda.synchronize()
