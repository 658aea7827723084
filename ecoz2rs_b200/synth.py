"""Synthetic LPC predictor frames: the `ar_mix` generator of SURVEY.md §8(d).

The reference ships no `.prd` data (its whale recordings are git-ignored), so
every benchmark and parity test runs on frames drawn from a mixture of
AR processes.  A frame is what `lpa_on_signal` would have produced for a
window of such a process: the exact autocorrelation `r[0..P]` (r0 = 1) divided
by the order-P prediction error energy (gain normalisation, reference
`ecoz2/src/lpc/lpa_on_signal.c:143-150`).

    sigma_m = 0.6 exp(-(m-1)/5) + 0.03               m = 1..P
    g_c     ~ N(0, sigma_m^2)                         C centres
    g       = g_c[label] + 0.3 sigma_m N(0,1)         label ~ U[0,C)
    k       = clip(tanh g, +-0.98)                    reflection coefficients
    inverse Levinson (r0 = 1, E0 = 1):
        r_m = -k_m E_{m-1} - sum_{i=1}^{m-1} a_i r_{m-i};  step-up a;  E_m = E_{m-1} (1 - k_m^2)
    frame   = r / E_P

Two back ends with the same formulae: NumPy (CPU tests, golden fixtures) and
torch (bulk generation on the GPU box for the 10M-frame benchmark shapes; only
the random streams differ).  This is input synthesis, not part of the VQ path.
"""
from __future__ import annotations

import numpy as np

DEFAULT_CENTRES = 256


def _sigma(P: int) -> np.ndarray:
    m = np.arange(1, P + 1, dtype=np.float64)
    return 0.6 * np.exp(-(m - 1.0) / 5.0) + 0.03


def _inverse_levinson(k, xp):
    """k: [n, P] reflection coefficients -> gain-normalised autocorrelation [n, P+1]."""
    n, P = k.shape
    if xp is np:
        r = np.zeros((n, P + 1), dtype=np.float64)
        a = np.zeros((n, P + 1), dtype=np.float64)
        E = np.ones(n, dtype=np.float64)
    else:
        r = xp.zeros((n, P + 1), dtype=xp.float64, device=k.device)
        a = xp.zeros((n, P + 1), dtype=xp.float64, device=k.device)
        E = xp.ones(n, dtype=xp.float64, device=k.device)
    r[:, 0] = 1.0
    a[:, 0] = 1.0
    for m in range(1, P + 1):
        km = k[:, m - 1]
        acc = -km * E
        if m > 1:
            # sum_{i=1}^{m-1} a_i r_{m-i}
            if xp is np:
                acc = acc - np.einsum("ij,ij->i", a[:, 1:m], r[:, m - 1:0:-1])
            else:
                acc = acc - (a[:, 1:m] * xp.flip(r[:, 1:m], dims=(1,))).sum(dim=1)
        r[:, m] = acc
        # step-up: a_i <- a_i + k_m a_{m-i}, a_m = k_m
        if m > 1:
            if xp is np:
                prev = a[:, 1:m].copy()
                a[:, 1:m] = prev + km[:, None] * prev[:, ::-1]
            else:
                prev = a[:, 1:m].clone()
                a[:, 1:m] = prev + km[:, None] * xp.flip(prev, dims=(1,))
        a[:, m] = km
        E = E * (1.0 - km * km)
    return r / E[:, None]


def ar_mix(N: int, P: int, C: int = DEFAULT_CENTRES, seed: int = 0,
           chunk: int = 1 << 18) -> np.ndarray:
    """NumPy back end.  Returns float64 [N, P+1]."""
    rng = np.random.default_rng(seed)
    sig = _sigma(P)
    centres = rng.normal(size=(C, P)) * sig
    out = np.empty((N, P + 1), dtype=np.float64)
    for lo in range(0, N, chunk):
        n = min(chunk, N - lo)
        labels = rng.integers(0, C, size=n)
        g = centres[labels] + 0.3 * sig * rng.normal(size=(n, P))
        k = np.clip(np.tanh(g), -0.98, 0.98)
        out[lo:lo + n] = _inverse_levinson(k, np)
    return out


def ar_mix_torch(N: int, P: int, C: int = DEFAULT_CENTRES, seed: int = 0,
                 device="cuda", chunk: int = 1 << 20, out=None, stream_id: int = 0):
    """torch back end (same formulae; torch's random stream).  float64 [N, P+1] on `device`.
    `stream_id` selects an independent frame stream of the SAME mixture (one per rank)."""
    import torch

    rng = np.random.default_rng(seed)
    sig_np = _sigma(P)
    centres = torch.from_numpy(rng.normal(size=(C, P)) * sig_np).to(device)
    sig = torch.from_numpy(sig_np).to(device)
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed) + 1000003 * int(stream_id))
    if out is None:
        out = torch.empty((N, P + 1), dtype=torch.float64, device=device)
    for lo in range(0, N, chunk):
        n = min(chunk, N - lo)
        labels = torch.randint(0, C, (n,), generator=gen, device=device)
        noise = torch.randn((n, P), generator=gen, device=device, dtype=torch.float64)
        g = centres[labels] + 0.3 * sig * noise
        k = torch.clamp(torch.tanh(g), -0.98, 0.98)
        out[lo:lo + n] = _inverse_levinson(k, torch)
    return out


def ar_mix_torch_range(lo: int, hi: int, P: int, C: int = DEFAULT_CENTRES, seed: int = 0, device="cuda",
                       chunk: int = 1 << 20):
    """Frames [lo, hi) of ONE global stream (torch back end): the stream is defined chunk by chunk
    (chunk c = frames [c * chunk, (c + 1) * chunk), its own generator seeded by (seed, c)), so any
    rank can produce exactly its slice and the union over the ranks of a multi-GPU run is the
    single-GPU data set.  float64 [hi - lo, P+1] on `device`."""
    import torch

    rng = np.random.default_rng(seed)
    sig_np = _sigma(P)
    centres = torch.from_numpy(rng.normal(size=(C, P)) * sig_np).to(device)
    sig = torch.from_numpy(sig_np).to(device)
    out = torch.empty((max(hi - lo, 0), P + 1), dtype=torch.float64, device=device)
    gen = torch.Generator(device=device)
    for c in range(lo // chunk, (max(hi, lo + 1) - 1) // chunk + 1):
        a, b = max(lo, c * chunk), min(hi, (c + 1) * chunk)
        if a >= b:
            continue
        gen.manual_seed(int(seed) * 1000003 + 7919 * c + 1)
        labels = torch.randint(0, C, (chunk,), generator=gen, device=device)
        noise = torch.randn((chunk, P), generator=gen, device=device, dtype=torch.float64)
        sl = slice(a - c * chunk, b - c * chunk)
        g = centres[labels[sl]] + 0.3 * sig * noise[sl]
        k = torch.clamp(torch.tanh(g), -0.98, 0.98)
        out[a - lo:b - lo] = _inverse_levinson(k, torch)
    return out
