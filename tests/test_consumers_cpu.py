"""CPU checks of the record consumers' restatement (tests/consumers.py) and of the AX.25 printer facade against the
unmodified reference (oracle/_ref)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import consumers
import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (needs /root/reference)")


@needs_ref
def test_prbs9_restatement_matches_reference_struct():
    rng = np.random.default_rng(5)
    gen = consumers.Prbs9()
    # a PRBS9 sequence with error bursts and resets in between
    bits = []
    st = 0x1A5
    for _ in range(6000):
        b = ((st >> 8) ^ (st >> 4)) & 1
        st = ((st << 1) | b) & 0x1FF
        bits.append(b)
    bits = np.array(bits, np.uint8)
    flips = rng.random(6000) < 0.02
    flips[2500:2700] = rng.random(200) < 0.4            # a burst that unlocks the synchroniser
    bits ^= flips.astype(np.uint8)
    resets = np.zeros(6000, np.uint8); resets[[1500, 4100]] = 1
    out = np.zeros(3, np.uint32)
    ol.ref().m17ref_prbs9_validate(bits.ctypes.data_as(ctypes.c_void_p), resets.ctypes.data_as(ctypes.c_void_p),
                                   ctypes.c_size_t(6000), out.ctypes.data_as(ctypes.c_void_p))
    for i in range(6000):
        if resets[i]:
            gen.reset()
        gen.validate(int(bits[i]))
    assert (int(gen.synced), gen.bit_count, gen.err_count) == tuple(int(v) for v in out)
    assert gen.err_count > 30 and gen.bit_count > 1000


def _ax25_address(call, ssid, last):
    b = bytearray((call + "      ")[:6].encode()) + bytearray([ssid & 0x0F])
    b = bytearray((c << 1) & 0xFF for c in b)
    b[6] |= 0x60 | (1 if last else 0)
    return bytes(b)


@needs_ref
def test_ax25_printer_matches_reference_output(tmp_path):
    exe = os.path.join(ROOT, "m17_cxx_demod_b200", "m17-demod")
    rng = np.random.default_rng(11)
    frames = []
    for k in range(40):
        n_rep = int(rng.integers(0, 4))
        f = _ax25_address("APRS", 0, False) + _ax25_address(f"N{k}CALL"[:6], k % 16, n_rep == 0)
        for r in range(n_rep):
            f += _ax25_address(f"WIDE{r + 1}", r + 1, r == n_rep - 1)
        f += bytes([(0x03, 0x00, 0x01, 0x13)[k % 4]])
        if k % 4 in (0, 3):
            f += b"\xF0"
        f += bytes(rng.integers(0, 256, int(rng.integers(0, 60)), dtype=np.uint8)) if k % 3 == 0 else f">status {k} test".encode()
        f += bytes(rng.integers(0, 256, 2, dtype=np.uint8))
        frames.append(f)
    frames += [b"", b"short", bytes(16), bytes(range(17)), bytes(rng.integers(0, 256, 300, dtype=np.uint8))]
    for i, f in enumerate(frames):
        p = tmp_path / f"f{i}.bin"
        p.write_bytes(f)
        # (through the reference binary rather than libm17ref.so: the printer uses iostreams, and the library's C++
        # runtime is not the one the Python process has loaded)
        want = subprocess.run([ol.ref_binary(), "--ax25-print", str(p)], capture_output=True, check=True).stdout
        got = subprocess.run([exe, "--ax25-print", str(p)], capture_output=True, check=True).stdout
        assert got == want, (i, f.hex())
