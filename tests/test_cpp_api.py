"""The reference's C++ API boundary (include/lora_phy/phy.hpp + liblora_phy.so) and its two drivers.

bin/rx_runner and bin/tx_runner are the reference's own runners/rx_runner.cpp and tx_runner.cpp, compiled
unmodified against this repo's headers and library (package Makefile).  On the GPU they must print / write exactly
what the unmodified reference does (goldens made by scripts/make_golden.py and make_golden_runners.py); without a
GPU they must fail loudly instead of computing anything on the CPU."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200")
G = os.path.join(ROOT, "tests", "golden")
RX, TX = os.path.join(PKG, "bin", "rx_runner"), os.path.join(PKG, "bin", "tx_runner")

# every function include/lora_phy/phy.hpp declares, as g++ mangles it (reference phy.hpp:102-156,190-215)
CPP_SYMBOLS = [
    "_ZN8lora_phy4initEPNS_14lora_workspaceEPKNS_11lora_paramsE",
    "_ZN8lora_phy5resetEPNS_14lora_workspaceE",
    "_ZN8lora_phy6encodeEPNS_14lora_workspaceEPKhmPtm",
    "_ZN8lora_phy6decodeEPNS_14lora_workspaceEPKtmPhm",
    "_ZN8lora_phy8modulateEPNS_14lora_workspaceEPKtmPSt7complexIfEm",
    "_ZN8lora_phy10demodulateEPNS_14lora_workspaceEPKSt7complexIfEmPtm",
    "_ZN8lora_phy16estimate_offsetsEPNS_14lora_workspaceEPKSt7complexIfEm",
    "_ZN8lora_phy18compensate_offsetsEPKNS_14lora_workspaceEPSt7complexIfEm",
    "_ZN8lora_phy16get_last_metricsEPKNS_14lora_workspaceE",
    "_ZN8lora_phy15lora_demod_initEPNS_20lora_demod_workspaceEjNS_11window_typeEPSt7complexIfEm",
    "_ZN8lora_phy15lora_demod_freeEPNS_20lora_demod_workspaceE",
    "_ZN8lora_phy13lora_modulateEPKtmPSt7complexIfEjjNS_9bandwidthEfh",
    "_ZN8lora_phy15lora_demodulateEPNS_20lora_demod_workspaceEPKSt7complexIfEmPtjPh",
    "_ZN8lora_phy11lora_encodeEPKhmPtj",
    "_ZN8lora_phy11lora_decodeEPKtmPh",
]


def test_cpp_library_exports_the_reference_api():
    lib = os.path.join(PKG, "liblora_phy.so")
    assert os.path.exists(lib), "run __graft_entry__.build()"
    names = subprocess.run(["nm", "-D", "--defined-only", lib], check=True, capture_output=True, text=True).stdout
    for sym in CPP_SYMBOLS:
        assert sym in names, sym


@pytest.mark.skipif(not os.path.exists(RX), reason="runners are built only where the reference tree is present")
def test_runners_fail_loudly_without_a_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    iq = np.zeros(256, np.complex64)
    f = tmp_path / "z.iq"
    iq.tofile(f)
    r = subprocess.run([RX, f"--in={f}", "--sf=7"], capture_output=True, text=True)
    assert r.returncode != 0 and "Payload" not in r.stdout
    r = subprocess.run([TX, "--payload=deadbeef", "--sf=7", f"--out={tmp_path / 'o.iq'}"], capture_output=True, text=True)
    assert r.returncode != 0


def test_stream_receiver_cli_fails_loudly_without_a_gpu(tmp_path):
    """bin/lora_rx_stream is built everywhere (it needs no reference source); with no device it must say so and
    print no frames -- there is no CPU path behind it."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    exe = os.path.join(os.path.dirname(RX), "lora_rx_stream")
    assert os.path.exists(exe), "bin/lora_rx_stream was not built"
    f = tmp_path / "z.iq"
    np.zeros(4096, np.complex64).tofile(f)
    r = subprocess.run([exe, f"--in={f}", "--sf=7"], capture_output=True, text=True)
    assert r.returncode != 0 and "Frames" not in r.stdout and "no CUDA device" in r.stderr
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "Usage" in r.stderr


needs_runners = pytest.mark.skipif(not os.path.exists(RX), reason="bin/rx_runner not built (no reference tree at build time)")


@pytest.mark.gpu
@needs_runners
def test_rx_runner_config1_golden_capture(tmp_path):
    """BASELINE configs[0]: stdout of the reference's rx_runner on the committed SF7 capture, byte for byte
    (sha256 recorded in SURVEY.md F7)."""
    g = np.load(os.path.join(G, "capture_sf7.npz"))
    f = tmp_path / "capture.iq"
    g["iq"].tofile(f)
    out = subprocess.run([RX, f"--in={f}", "--sf=7", "--cr=1", "--bw=125000", "--report-offsets"], check=True,
                         capture_output=True).stdout
    assert out == g["rx_runner_stdout"].tobytes()
    assert hashlib.sha256(out).hexdigest() == "91e55ba04aaff03239abd76229b6a6b13ec139f2e351439f92f87f726ac2cddb"


@pytest.mark.gpu
@needs_runners
def test_tx_rx_runner_loopback_matches_reference(tmp_path):
    """tx_runner writes the reference's IQ bit for bit, and rx_runner prints what the reference prints for it --
    including the reference's own non-round-trip through demodulate()."""
    g = np.load(os.path.join(G, "runners_loopback.npz"))
    for k in range(3):
        payload, sf, bw = (str(x) for x in g[f"case{k}_args"])
        f = tmp_path / f"tx{k}.iq"
        subprocess.run([TX, f"--payload={payload}", f"--sf={sf}", f"--bw={bw}", f"--out={f}"], check=True)
        assert np.fromfile(f, np.complex64).tobytes() == g[f"case{k}_iq"].tobytes(), (k, "tx IQ differs")
        out = subprocess.run([RX, f"--in={f}", f"--sf={sf}", f"--bw={bw}", "--report-offsets"], check=True,
                             capture_output=True).stdout
        assert out == g[f"case{k}_rx_stdout"].tobytes(), (k, out)


PROFILES_YAML = """-
  name: sf7_bw125_cr45
  sf: 7
  bw: 125000
  cr: 4/5
-
  name: sf7_bw125_cr47
  sf: 7
  bw: 125000
  cr: 4/7
-
  name: sf8_bw125_cr45
  sf: 8
  bw: 125000
  cr: 4/5
"""


@pytest.mark.gpu
@needs_runners
@pytest.mark.parametrize("name", ["e2e_chain_test", "bit_exact_test", "roundtrip_test", "equal_power_bin_test",
                                  "gr_lora_sdr_interop"])
def test_reference_test_programs_pass_against_this_library(tmp_path, name):
    """The reference's own tests/<name>.cpp, compiled unmodified against include/lora_phy and liblora_phy.so
    (SURVEY.md section 4): encode -> lora_modulate -> caller dechirp -> lora_demodulate -> decode round trips for
    the three profiles, the DEADBEEF known answers, the equal-power tie rule and the osr=2 interop capture."""
    exe = os.path.join(PKG, "bin", "ref_" + name)
    if not os.path.exists(exe):
        pytest.skip("not built")
    (tmp_path / "tests").mkdir()
    (tmp_path / "tests" / "profiles.yaml").write_text(PROFILES_YAML)   # the three profiles of tests/profiles.yaml
    np.load(os.path.join(G, "interop_sf7_osr2.npz"))["iq"].tofile(tmp_path / "test_output.iq")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, (name, r.stdout[-400:], r.stderr[-400:])


@needs_runners
@pytest.mark.parametrize("name", ["whitening_test", "hamming_cross_check"])
def test_reference_code_helper_tests_pass_against_this_header(name):
    """tests/whitening_test.cpp and tests/hamming_cross_check.cpp of the reference, unmodified, against this build's
    include/lora_phy/LoRaCodes.hpp (the second one compares it with the upstream LoRa-SDR header bit flip by bit flip).
    Caller-side helpers: they need no device."""
    exe = os.path.join(PKG, "bin", "ref_" + name)
    if not os.path.exists(exe):
        pytest.skip("not built")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, (name, r.stdout[-400:], r.stderr[-400:])


REF_DUMP = os.path.join(ROOT, "oracle", "_ref", "lora_phy_vector_dump")
OUR_DUMP = os.path.join(PKG, "bin", "lora_phy_vector_dump")
DUMP_FILES = ("payload.bin", "pre_interleave.csv", "post_interleave.csv", "iq_samples.csv", "demod_symbols.csv",
              "deinterleave.csv", "decoded.bin")


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(REF_DUMP) and os.path.exists(OUR_DUMP)), reason="vector dump binaries not built")
@pytest.mark.parametrize("sf,extra", [(7, []), (8, []), (7, ["--window=hann"]), (9, ["--bytes=32", "--bw=250000"]),
                                      (12, ["--bytes=4"])])
def test_vector_dump_stage_files_equal_the_reference(tmp_path, sf, extra):
    """runners/lora_phy_vector_dump.cpp compiled unmodified against this build
    (encode -> modulate -> demodulate on the GPU, the code helpers from this build's LoRaCodes.hpp) writes the seven
    stage dumps byte for byte as the reference binary does, for the profiles of tests/profiles.yaml and seeds 0..2;
    scripts/generate_vectors.py then produces the same manifest (sha256 per base64-encoded file) from either."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import generate_vectors
    for seed in range(3):
        outs = {}
        for who, exe in (("ref", REF_DUMP), ("b200", OUR_DUMP)):
            d = tmp_path / f"{who}_{seed}"
            d.mkdir()
            r = subprocess.run([exe, f"--sf={sf}", f"--seed={seed}", f"--out={d}"] + extra, capture_output=True, text=True)
            assert r.returncode == 0, (who, r.stderr[-300:])
            outs[who] = d
        for name in DUMP_FILES:
            a, b = (outs["ref"] / name).read_bytes(), (outs["b200"] / name).read_bytes()
            assert len(a) > 0 and a == b, (sf, seed, name)
        man = {}
        for who, exe in (("ref", REF_DUMP), ("b200", OUR_DUMP)):
            args = ["--sf", str(sf), "--seed", str(seed), "--out", f"m_{who}_{seed}", "--root", str(tmp_path), "--binary", exe,
                    "--cfo-bins", "0.25", "--time-offset", "3"]
            for e in extra:
                k, v = e.split("=")
                args += [k, v]
            man[who] = generate_vectors.main(args)
        assert man["ref"] == man["b200"] and len(man["ref"]["files"]) == 8


@pytest.mark.gpu
@needs_runners
def test_no_alloc_test_report(tmp_path):
    """tests/no_alloc_test.cpp of the reference, unmodified, must pass against the drop-in (zero allocations seen in
    lora_modulate / lora_demod_init / lora_demodulate, symbols [0,1,12,34,56] round-trip exactly).  Its tracker is an
    `inline operator new` in a header, so it sees the allocations of the translation unit it is compiled into; the
    stricter bin/no_alloc_steady exports a global operator new that also counts what liblora_phy.so,
    liblora_phy_b200.so and the CUDA runtime inside them allocate: there the first call for a spreading factor
    creates the shared device context (what the reference's init does inside the caller's workspace; measured: 9
    allocations in the first lora_modulate, 1 in the first lora_demodulate) and every later call must allocate
    nothing."""
    exe = os.path.join(PKG, "bin", "ref_no_alloc_test")
    if os.path.exists(exe):
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode == 0 and "No allocations detected" in r.stdout, (r.stdout, r.stderr)
    steady = os.path.join(PKG, "bin", "no_alloc_steady")
    assert os.path.exists(steady), "bin/no_alloc_steady was not built"
    r = subprocess.run([steady], capture_output=True, text=True)
    assert r.returncode == 0 and "No allocations detected" in r.stdout, (r.stdout, r.stderr)
