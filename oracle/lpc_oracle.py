"""CPU restatement of the reference's LPC front end (signal -> gain-normalised
autocorrelation vectors, the `.prd` content the VQ path consumes).

TEST INFRASTRUCTURE ONLY: imported by tests/, `__graft_entry__.smoke()` and
bench.py's cpu_baseline leg, never by the product.

Follows, line by line:
  * `lpa_on_signal`  /root/reference/ecoz2/src/lpc/lpa_on_signal.c:81-157
    (frame count :96-101, Hamming :24-31, mean removal :42-57, pre-emphasis
    :59-71, window :73-80, gain normalisation :143-150)
  * `lpca`           /root/reference/ecoz2/src/lpc/lpca.c:9-59
    (autocorrelation :16-21, Levinson-Durbin recursion :27-47)

Pinned against the reference itself (oracle/_ref/<arch>/libecoz2_ref_lpc.so,
compiled from those files) on the reference's own test fixture
`signal_frame.inputs` (src/lpc/lpca_rs.rs:214-306) and on seeded signals:
tests/golden/make_golden_lpc.py, tests/test_lpc_cpu.py.
All arithmetic is FP64, vectorised over frames; sums run in a different order
from the C loops, so agreement is to rounding (checked at 1e-9 relative, the
tolerance the reference's own test uses between its lpca variants).
"""
from __future__ import annotations

import numpy as np


def frame_geometry(num_samples: int, sample_rate: float, window_ms: int, offset_ms: int):
    """-> (winSize, offset, T) exactly as lpa_on_signal.c:84-101 computes them
    (sampleRate truncated to long, integer division)."""
    sr = int(sample_rate)
    win = int((window_ms * sr) // 1000)
    off = int((offset_ms * sr) // 1000)
    if win > num_samples:
        return win, off, -1          # "signal too short"
    if off <= 0:
        raise ValueError("offset of zero samples")
    t = int((num_samples - (win - off)) // off)
    if (t - 1) * off + win > num_samples:
        t -= 1
    return win, off, t


def hamming(win: int) -> np.ndarray:
    n = np.arange(win, dtype=np.float64)
    return .54 - .46 * np.cos((n * 2 * np.pi) / (win - 1))


def windowed_frames(signal: np.ndarray, win: int, off: int, T: int) -> np.ndarray:
    """[T, win] frames after mean removal, pre-emphasis (0.95) and Hamming window."""
    x = np.ascontiguousarray(signal, dtype=np.float64)
    fr = np.lib.stride_tricks.sliding_window_view(x, win)[::off][:T].copy()
    fr -= fr.sum(axis=1, keepdims=True) / win
    pre = fr.copy()
    pre[:, 1:] -= .95 * fr[:, :-1]          # x[n] -= .95 x[n-1], done from the end backwards
    return pre * hamming(win)[None, :]


def lpca(frames: np.ndarray, P: int):
    """lpca (lpca.c:9-59) on every row of frames.
    -> r [T, P+1], rc [T, P+1], a [T, P+1], pe [T], status [T] (0 ok, 1 r0 == 0, 2 pe <= 0)."""
    T, N = frames.shape
    r = np.empty((T, P + 1))
    for i in range(P + 1):
        r[:, i] = np.einsum("tn,tn->t", frames[:, :N - i], frames[:, i:])
    status = np.zeros(T, dtype=np.int32)
    status[r[:, 0] == 0.] = 1
    pe = r[:, 0].copy()
    a = np.zeros((T, P + 1))
    rc = np.zeros((T, P + 1))
    a[:, 0] = 1.
    live = status == 0
    with np.errstate(all="ignore"):
        for k in range(1, P + 1):
            s = np.zeros(T)
            for i in range(1, k + 1):
                s -= a[:, k - i] * r[:, i]
            akk = s / pe
            upd = live.copy()
            rc[upd, k] = akk[upd]
            new = a.copy()
            new[:, k] = akk
            for i in range(1, (k >> 1) + 1):
                ai, aj = a[:, i], a[:, k - i]
                new[:, i] = ai + akk * aj
                new[:, k - i] = aj + akk * ai
            a[upd] = new[upd]
            pe2 = pe * (1. - akk * akk)
            pe[upd] = pe2[upd]
            dead = upd & (pe <= 0.)
            status[dead] = 2
            live &= ~dead
    return r, rc, a, pe, status


def lpa_on_signal(signal: np.ndarray, sample_rate: float, P: int, window_ms: int, offset_ms: int):
    """-> vectors [T, P+1] = r / errPred per frame (left unnormalised where errPred == 0),
    or None when the signal is shorter than one window."""
    win, off, T = frame_geometry(len(signal), sample_rate, window_ms, offset_ms)
    if T < 0:
        return None
    if T == 0:
        return np.zeros((0, P + 1))
    fr = windowed_frames(signal, win, off, T)
    r, _, _, pe, status = lpca(fr, P)
    out = r.copy()
    # lpca returns early on errors: status 1 leaves pe unset (the reference then divides by
    # whatever errPred holds -- undefined; such all-zero frames are excluded from parity),
    # status 2 leaves pe <= 0 and the vector is still divided by it when non-zero
    nz = (pe != 0.) & (status != 1)
    out[nz] /= pe[nz, None]
    return out
