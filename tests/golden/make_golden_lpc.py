"""Generates tests/golden/lpc_front.npz: golden vectors of the LPC front end.

Run in the build container (needs /root/reference and oracle/_ref):
    python tests/golden/make_golden_lpc.py

Contents
  frame_x                the reference's own test fixture `signal_frame.inputs` (CBOR
                         {x: [f64], p: uint}; src/lpc/lpca_rs.rs:214-306 runs lpca on it at P = 36)
  frame_{r,rc,a,pe}      the reference C `lpca` (ecoz2/src/lpc/lpca.c:9-59) on that frame
  sig_*                  a seeded synthetic signal and the reference `lpa_on_signal`
                         (ecoz2/src/lpc/lpa_on_signal.c:81-157) vectors for several
                         (P, window, offset) settings
"""
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbuild  # noqa: E402

FIXTURE = "/root/reference/signal_frame.inputs"


def cbor_lpca_input(path):
    """Minimal CBOR reader for {"x": [float64...], "p": uint} (serde_cbor output)."""
    b = open(path, "rb").read()
    pos = 0

    def head():
        nonlocal pos
        ib = b[pos]; pos += 1
        major, info = ib >> 5, ib & 31
        if info < 24:
            val = info
        elif info == 24:
            val = b[pos]; pos += 1
        elif info == 25:
            val = struct.unpack(">H", b[pos:pos + 2])[0]; pos += 2
        elif info == 26:
            val = struct.unpack(">I", b[pos:pos + 4])[0]; pos += 4
        elif info == 27:
            val = struct.unpack(">Q", b[pos:pos + 8])[0]; pos += 8
        else:
            raise ValueError("indefinite length not expected")
        return major, info, val

    def item():
        nonlocal pos
        major, info, val = head()
        if major == 0:
            return val
        if major == 3:
            s = b[pos:pos + val].decode(); pos += val
            return s
        if major == 4:
            return [item() for _ in range(val)]
        if major == 5:
            return {item(): item() for _ in range(val)}
        if major == 7 and info == 27:
            return struct.unpack(">d", struct.pack(">Q", val))[0]
        if major == 7 and info == 26:
            return struct.unpack(">f", struct.pack(">I", val))[0]
        raise ValueError(f"unexpected CBOR item {major}/{info}")

    d = item()
    return np.array(d["x"], dtype=np.float64), int(d["p"])


def synth_signal(n, seed, sample_rate=32000.0):
    """Seeded test signal: a few drifting tones + coloured noise + a silent stretch."""
    rng = np.random.default_rng(seed)
    t = np.arange(n) / sample_rate
    x = np.zeros(n)
    for f0, am in ((310.0, 0.30), (1270.0, 0.18), (3650.0, 0.07)):
        x += am * np.sin(2 * np.pi * (f0 * t + 40.0 * np.sin(2 * np.pi * 0.7 * t)))
    w = rng.standard_normal(n)
    w[1:] += 0.7 * w[:-1]
    x += 0.05 * w
    x += 0.02                                # DC offset (exercises mean removal)
    x = np.round(x * 32767.0 * 0.5) / 32768.0   # 16-bit PCM grid, as a wav file delivers
    return x.astype(np.float32).astype(np.float64)


def main():
    assert refbuild.lpc_available(), "build oracle/_ref first (make -C oracle ref)"
    out = {}
    x, p = cbor_lpca_input(FIXTURE)
    print(f"fixture: {len(x)} samples, p = {p}")
    st, r, rc, a, pe = refbuild.ref_lpca(x, 36)
    assert st == 0
    out.update(frame_x=x, frame_p=np.int32(p), frame_r=r, frame_rc=rc, frame_a=a, frame_pe=np.float64(pe))
    sig = synth_signal(48000, 20260104)
    out["sig_x"] = sig
    out["sig_rate"] = np.float64(32000.0)
    cfgs = [(36, 45, 15), (12, 45, 15), (20, 30, 10), (36, 25, 25), (44, 45, 15)]
    out["sig_cfgs"] = np.array(cfgs, dtype=np.int32)
    for i, (P, W, O) in enumerate(cfgs):
        v = refbuild.ref_lpa_on_signal(sig, 32000.0, P, W, O, threads=1)
        print(f"cfg {i}: P={P} W={W} O={O} -> {v.shape}")
        out[f"sig_v{i}"] = v
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "lpc_front.npz"), **out)
    print("wrote tests/golden/lpc_front.npz")


if __name__ == "__main__":
    main()
