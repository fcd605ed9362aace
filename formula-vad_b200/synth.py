"""Seeded synthetic onboard-like audio: engine-harmonic noise with slow RPM sweeps
plus broadband noise, and speech-like bursts (formant-band AM carriers with a syllabic envelope).
Returns planar float32 PCM and the ground-truth burst labels (seconds)."""
from __future__ import annotations

import numpy as np


def onboard_stream(seed: int, seconds: float, sample_rate: int = 16000, n_channels: int = 1,
                   bursts_per_10min: tuple[int, int] = (8, 40)):
    rng = np.random.default_rng(seed)
    n = int(round(seconds * sample_rate))
    t = np.arange(n, dtype=np.float64) / sample_rate
    # engine fundamental: slow random walk in 60..300 Hz (piecewise-linear between knots every 2 s)
    n_knots = max(2, int(seconds / 2) + 2)
    walk = np.clip(150 + np.cumsum(rng.normal(0, 18, n_knots)), 60, 300)
    f0 = np.interp(t, np.linspace(0, seconds, n_knots), walk)
    phase = 2 * np.pi * np.cumsum(f0) / sample_rate
    engine = np.zeros(n)
    for h in range(1, 13):
        if h * 300 < sample_rate / 2:
            engine += (1.0 / h) * np.sin(h * phase + rng.uniform(0, 2 * np.pi))
    engine *= 0.05
    # speech-like bursts
    speech = np.zeros(n)
    labels = []
    per10 = rng.integers(bursts_per_10min[0], bursts_per_10min[1] + 1)
    n_bursts = max(1, int(round(per10 * seconds / 600.0)))
    cursor = rng.uniform(1.0, 4.0)
    for _ in range(n_bursts):
        dur = rng.uniform(0.8, 12.0)
        if cursor + dur >= seconds - 0.5:
            break
        i0, i1 = int(cursor * sample_rate), int((cursor + dur) * sample_rate)
        tt = t[i0:i1] - t[i0]
        env = 0.55 + 0.45 * np.sin(2 * np.pi * 4.0 * tt + rng.uniform(0, 2 * np.pi))
        edge = np.minimum(1.0, np.minimum(tt, tt[-1] - tt) / 0.05)
        sig = np.zeros(i1 - i0)
        for _k in range(rng.integers(3, 6)):
            fc = rng.uniform(300, 3400)
            sig += np.sin(2 * np.pi * fc * tt + rng.uniform(0, 2 * np.pi)) * rng.uniform(0.5, 1.0)
        amp = 10 ** (rng.uniform(6, 15) / 20) * 0.04
        speech[i0:i1] += amp * env * edge * sig
        labels.append((cursor, cursor + dur))
        cursor += dur + rng.uniform(3.0, 3.0 + 2 * 600.0 / max(per10, 1))
    chans = []
    for c in range(n_channels):
        scale = 1.0 if c == 0 else rng.uniform(0.3, 1.0)
        noise = rng.normal(0, 10 ** (-30 / 20) * 0.3, n)
        x = scale * engine + noise + speech
        chans.append(np.clip(x, -1, 1).astype(np.float32))
    return np.stack(chans), labels


def pcm16(x: np.ndarray) -> np.ndarray:
    """float [-1, 1) -> int16 the way a 16-bit WAV stores it (round to nearest, clipped)."""
    return np.clip(np.rint(np.asarray(x, np.float64) * 32768.0), -32768, 32767).astype(np.int16)


def expand_16k_to_48k(x16: np.ndarray) -> np.ndarray:
    """x48[3i] = x16[i], the two samples in between linearly interpolated: the
    reference's decimate-by-3 then recovers x16 exactly."""
    x16 = np.asarray(x16, np.float32)
    n = x16.shape[-1]
    out = np.empty(x16.shape[:-1] + (3 * n,), np.float32)
    nxt = np.concatenate([x16[..., 1:], x16[..., -1:]], axis=-1)
    out[..., 0::3] = x16
    out[..., 1::3] = x16 + (nxt - x16) * np.float32(1 / 3)
    out[..., 2::3] = x16 + (nxt - x16) * np.float32(2 / 3)
    return out


def onboard_batch_torch(seeds, seconds: float, sample_rate: int = 16000, device="cuda"):
    """Device-side generator of the same recipe for benchmark-sized batches (hours x streams):
    returns a float32 tensor [len(seeds), n] on `device` and the per-stream burst labels (seconds).
    torch is plumbing here (device memory + elementwise math for synthetic data), not the product."""
    import torch

    n = int(round(seconds * sample_rate))
    out = torch.empty((len(seeds), n), dtype=torch.float32, device=device)
    all_labels = []
    piece = 1 << 24
    for si, seed in enumerate(seeds):
        rng = np.random.default_rng(seed)
        n_knots = max(2, int(seconds / 2) + 2)
        walk = np.clip(150 + np.cumsum(rng.normal(0, 18, n_knots)), 60, 300)
        knots = torch.tensor(walk, dtype=torch.float64, device=device)
        ph0 = rng.uniform(0, 2 * np.pi, 12)
        g = torch.Generator(device=device)
        g.manual_seed(int(seed) * 7919 + 17)
        phase_carry = 0.0
        for a in range(0, n, piece):
            b = min(n, a + piece)
            t = torch.arange(a, b, dtype=torch.float64, device=device) / sample_rate
            pos = t / seconds * (n_knots - 1)
            i0 = pos.floor().clamp_(0, n_knots - 2).long()
            fr = pos - i0
            f0 = knots[i0] * (1 - fr) + knots[i0 + 1] * fr
            ph = torch.cumsum(f0, 0) * (2 * np.pi / sample_rate) + phase_carry
            phase_carry = float(ph[-1] % (2 * np.pi))
            ph = (ph % (2 * np.pi)).float()
            eng = torch.zeros(b - a, dtype=torch.float32, device=device)
            for h in range(1, 13):
                if h * 300 < sample_rate / 2:
                    eng += (1.0 / h) * torch.sin(h * ph + float(ph0[h - 1]))
            noise = torch.randn(b - a, generator=g, device=device, dtype=torch.float32) * (10 ** (-30 / 20) * 0.3)
            out[si, a:b] = 0.05 * eng + noise
        # speech-like bursts
        labels = []
        per10 = rng.integers(8, 41)
        n_bursts = max(1, int(round(per10 * seconds / 600.0)))
        cursor = rng.uniform(1.0, 4.0)
        for _ in range(n_bursts):
            dur = rng.uniform(0.8, 12.0)
            if cursor + dur >= seconds - 0.5:
                break
            i0, i1 = int(cursor * sample_rate), int((cursor + dur) * sample_rate)
            tt = torch.arange(i1 - i0, dtype=torch.float32, device=device) / sample_rate
            env = 0.55 + 0.45 * torch.sin(2 * np.pi * 4.0 * tt + float(rng.uniform(0, 2 * np.pi)))
            edge = torch.minimum(tt, tt[-1] - tt).div_(0.05).clamp_(max=1.0)
            sig = torch.zeros_like(tt)
            for _k in range(rng.integers(3, 6)):
                fc = rng.uniform(300, 3400)
                sig += torch.sin(2 * np.pi * fc * tt + float(rng.uniform(0, 2 * np.pi))) * float(rng.uniform(0.5, 1.0))
            amp = 10 ** (rng.uniform(6, 15) / 20) * 0.04
            out[si, i0:i1] += amp * env * edge * sig
            labels.append((cursor, cursor + dur))
            cursor += dur + rng.uniform(3.0, 3.0 + 2 * 600.0 / max(per10, 1))
        all_labels.append(labels)
    out.clamp_(-1, 1)
    return out, all_labels
