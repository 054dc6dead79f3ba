"""Frequency-set tooling in front of the hot path: `Compressor` picks the <= max_size
frequency points of a measured FRF that the loss is evaluated on (it sets F, the number of
systems per candidate).  Mirrors `/root/reference/source/jax_plate/Input.py:5-158`: same
constructor, same call, and -- index selection being integer work -- the SAME selected
indices, including the reference's edge behaviour (alg 0: the running float position is
accumulated, not multiplied; alg 1: the fill-up loop advances a window edge by one and marks
the sample one past it, and counts a step even when that sample was already selected).
`tests/test_host.py` compares the selections with those of the reference's class.
"""
from __future__ import annotations

import numpy as np
from scipy.signal import find_peaks, peak_prominences, peak_widths, savgol_filter

UNIFORM, PEAKS = 0, 1


def _uniform_mask(size: int, want: int) -> np.ndarray:
    """alg 0 (`Input.py:71-81`): samples at int(0), int(step), int(step + step), ... with the
    position accumulated in floating point."""
    step = size / want
    # positions 0, step, 2 steps ... as the running sum the reference forms (cumsum adds
    # sequentially, so the rounding is the same); a few spare terms cover the loop's tail
    pos = np.concatenate(([0.0], np.cumsum(np.full(want + 8, step))))
    pos = pos[pos < size]
    mask = np.zeros(size, dtype=bool)
    mask[pos.astype(np.int64)] = True
    if mask.sum() > want:
        mask[0] = False
    return mask


def peak_smoothness(x: np.ndarray, peaks: np.ndarray) -> np.ndarray:
    """`Input.py:39-64`: inverse total variation of the (up to) 21 samples around each peak."""
    out = np.zeros(peaks.size, dtype=np.float64)
    for i, p in enumerate(peaks):
        half = 10 if (p > 10 and x.size - p > 10) else min(p, x.size - p) - 1
        seg = x[p - half:p + half + 1]
        out[i] = np.abs(np.diff(seg)).sum() / 2 / half * 20
    return 1.0 / out


def characteristic_points(freqs: np.ndarray, amplitude: np.ndarray) -> np.ndarray:
    """Resonances and anti-resonances kept by alg 1 (`Input.py:84-116`): extrema of the
    smoothed log-amplitude at least ~75 Hz apart, wider than 20 samples, with prominence > 0.1
    and smoothness < 50.  Maxima first, then minima, each ascending."""
    spacing = int(75 / np.max(np.diff(freqs)))
    log_afc = np.log(savgol_filter(np.abs(amplitude), 30, 3))
    found = []
    for sig in (log_afc, -log_afc):
        cand = find_peaks(sig, distance=spacing)[0]
        cand = cand[peak_widths(sig, cand)[0] > 20]
        cand = cand[peak_prominences(sig, cand)[0] > 0.1]
        found.append(cand[peak_smoothness(sig, cand) < 50])
    return np.concatenate(found)


def _peak_mask(size: int, want: int, centres: np.ndarray) -> np.ndarray:
    """alg 1 (`Input.py:118-156`): equal windows around every characteristic point, then the
    windows grow to the right one sample at a time until `want` steps have been spent."""
    n = centres.size
    half = (want - n) // (n * 2)
    lo, hi = centres - half, centres + half
    lo[lo < 0] = 0
    hi[hi + 1 > size] = size
    mask = np.zeros(size, dtype=bool)
    for a, b in zip(lo, hi):
        mask[a:b + 1] = True
    todo = want - int(mask.sum())
    while todo != 0:
        for i in range(n - 1):
            if hi[i] < lo[i + 1]:
                hi[i] += 1
                todo -= 1
                mask[hi[i] + 1] = True      # the reference marks the sample PAST the new edge
            if todo == 0:
                break
        if todo == 0:
            break
        if hi[-1] + 1 < size:
            hi[-1] += 1
            todo -= 1
            mask[hi[-1]] = True
        elif lo[0] - 1 > 0:
            lo[0] -= 1
            todo -= 1
            mask[lo[0]] = True
    return mask


class Compressor:
    """`Compressor(freqs, complex_fr, max_size, use_alg)(desired_size) -> (freqs, fr)`."""

    def __init__(self, freqs: np.ndarray, complex_fr: np.ndarray, max_size: int, use_alg: int):
        assert freqs.size == complex_fr.size
        self.size = freqs.size
        self.freqs, self.complex_fr = freqs, complex_fr
        self.max_size, self.alg = max_size, use_alg

    def mask(self, desired_size: int) -> np.ndarray:
        if desired_size > self.max_size:
            raise ValueError(f"Desired size of compressed data must be lower than {self.max_size + 1}")
        if self.alg == UNIFORM:
            return _uniform_mask(self.size, desired_size)
        if self.alg == PEAKS:
            return _peak_mask(self.size, desired_size, characteristic_points(self.freqs, self.complex_fr))
        return np.zeros(self.size, dtype=bool)      # the reference leaves other values unimplemented

    def __call__(self, desired_size: int):
        m = self.mask(desired_size)
        return self.freqs[m], self.complex_fr[m]
