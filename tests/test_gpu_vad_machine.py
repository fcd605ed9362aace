"""The VAD machine operator (fvad_vad_machine_scan) against the oracle's VADMachine (oracle_vad.c, which follows
src/AudioPipeline/VADMachine.zig:138-325 frame by frame) on prepared series: decisions must be identical and the
segments equal.  The series are built to reach every path of the run-based scan in vadscan.cu: long runs (one
warp prefix per 32 frames), a decision that flips every frame (the frame-by-frame fallback), the warm-up before
the long ring is full, a pre-filled ring, and state carried across launches."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SR, FFT = 48000, 1024


def _oracle_scan(po, cfg, band, ratio):
    L = po.lib()
    h = L.orc_vadm_new(C.byref(cfg), SR, 1, FFT)
    met = np.zeros(band.shape[0], np.uint8)
    sn = C.c_uint64(0)
    one = (C.c_float * 1)()
    try:
        for f in range(band.shape[0]):
            one[0] = float(band[f])
            L.orc_vadm_run(h, one, f * FFT, 1, float(ratio[f]), C.byref(sn))
            met[f] = L.orc_vadm_last_met(h)
        ptr = C.POINTER(po.Segment)()
        n = L.orc_vadm_segments(h, C.byref(ptr))
        segs = [(ptr[i].sample_from, ptr[i].sample_to, ptr[i].avg_channel_vol_ratio, ptr[i].vad_met_sec) for i in range(n)]
    finally:
        L.orc_vadm_free(h)
    return met, segs


def _series(kind, n, seed):
    rng = np.random.default_rng(seed)
    noise = (0.004 * np.exp(0.5 * rng.standard_normal(n))).astype(np.float32)
    ratio = np.clip(0.75 + 0.2 * rng.standard_normal(n), 0.0, 1.0).astype(np.float32)
    band = noise.copy()
    if kind == "bursts":       # speech in runs of 0.5-6 s, silence in between
        f = 300
        while f < n - 400:
            ln = int(rng.integers(25, 280))
            band[f:f + ln] *= np.float32(rng.uniform(30, 400))
            f += ln + int(rng.integers(40, 900))
    elif kind == "flicker":    # every other frame far above / far below the threshold
        band[2001:12000:2] *= np.float32(200.0)
        band[15001:15400:2] *= np.float32(200.0)
    elif kind == "edge":       # hovering around the threshold: short runs of random length
        band *= np.exp(2.0 * rng.standard_normal(n)).astype(np.float32) * np.float32(8.0)
    elif kind == "ratio":      # band always loud, the channel-ratio condition decides
        band *= np.float32(100.0)
        ratio = np.where(rng.random(n) < 0.5, 0.9, 0.1).astype(np.float32)
        ratio = np.repeat(ratio[::37], 37)[:n].astype(np.float32)
    return band, ratio


CASES = [
    ("bursts", dict(), 0),
    ("bursts", dict(has_initial_long_term_avg=0, long_term_speech_avg_sec=15.0), 0),     # warm-up, then prefix path
    ("flicker", dict(short_term_speech_avg_sec=0.03), 0),                               # n_short = 1: flips per frame
    ("edge", dict(short_term_speech_avg_sec=0.05, long_term_speech_avg_sec=20.0), 0),
    ("ratio", dict(channel_vol_ratio_avg_sec=0.1), 0),
    ("bursts", dict(speech_threshold_factor=4.0, max_speech_gap_sec=0.5, min_vad_duration_sec=0.2), 777),  # carried state
    ("edge", dict(long_term_speech_avg_sec=12.0, has_initial_long_term_avg=0), 1000),
]


@pytest.mark.parametrize("kind,over,step", CASES)
def test_vad_machine_scan_matches_oracle(fv, po, kind, over, step):
    n = 20000
    band, ratio = _series(kind, n, 11)
    g_met, g_segs = fv.VADMachine(fv.vad_machine_config(**over), SR, FFT).scan(band[None, :], ratio, frames_per_launch=step)
    o_met, o_segs = _oracle_scan(po, po.VadMachineConfig.default(**over), band, ratio)
    assert o_met.sum() > 0 and o_met.sum() < n                      # the series exercises both decisions
    assert np.array_equal(g_met[0], o_met), int((g_met[0] != o_met).sum())
    assert len(g_segs[0]) == len(o_segs)
    for g, o in zip(g_segs[0], o_segs):
        assert g[0] == o[0] and g[1] == o[1]
        assert abs(g[2] - o[2]) <= 1e-6 * max(1.0, abs(o[2])) and abs(g[3] - o[3]) <= 1e-5 * max(1.0, abs(o[3]))


def test_many_machines_in_one_launch(fv, po):
    """Every machine is an independent warp reading its own band row (BASELINE config 5's sweep axis)."""
    n = 12000
    overs = [dict(speech_threshold_factor=float(t), min_consecutive_sec_to_open=float(o)) for t in (3, 6, 10, 20) for o in (0.1, 0.2, 0.4)]
    rows, ratio = [], None
    for i in range(len(overs)):
        b, r = _series("bursts" if i % 2 == 0 else "edge", n, 100 + i)
        rows.append(b)
        ratio = r if ratio is None else ratio
    band = np.stack(rows)
    g_met, g_segs = fv.VADMachine([fv.vad_machine_config(**o) for o in overs], SR, FFT).scan(band, ratio)
    for m, o in enumerate(overs):
        o_met, o_segs = _oracle_scan(po, po.VadMachineConfig.default(**o), band[m], ratio)
        assert np.array_equal(g_met[m], o_met), m
        assert [(a, b) for a, b, _, _ in g_segs[m]] == [(a, b) for a, b, _, _ in o_segs], m


def test_segment_overflow_is_counted_not_written(fv):
    n = 6000
    band, ratio = _series("bursts", n, 5)
    vm = fv.VADMachine(fv.vad_machine_config(min_vad_duration_sec=0.0, max_speech_gap_sec=0.1), SR, FFT, max_segments=2)
    met, segs = vm.scan(band[None, :], ratio)
    assert len(segs[0]) == 2


@pytest.mark.parametrize("over,step", [
    (dict(), 0),                                     # warp kernel (run-based long-term recurrence), one launch
    (dict(), 37888),                                 # ... state carried across launches the size of an engine span
    (dict(long_term_speech_avg_sec=9.0), 0),         # n_long = 421 < 514: the one-thread-per-machine kernel
])
def test_multi_hour_loud_then_silent_series(fv, po, over, step):
    """Three hours whose first minutes are 10^7 times louder than the rest: while loud values share the long-term
    window with quiet ones the f64 running sum loses the quiet values' low bits, and once the loud ones are evicted a
    never-resynchronised sum would keep that error for the rest of the stream (the reference re-sums the ring on every
    push, RollingAverage.zig:45-56).  vadscan.cu re-derives the sum from the ring once per turn; the decisions must stay
    identical to the oracle's over the whole series."""
    n = int(3 * 3600 * SR / FFT)
    rng = np.random.default_rng(2024)
    band = (2e-7 * np.exp(0.4 * rng.standard_normal(n))).astype(np.float32)
    band[:30000] *= np.float32(1e7)                   # ~10 minutes around 2.0
    ratio = np.full(n, 0.9, np.float32)
    f = 35000
    while f < n - 2000:                                # speech bursts 8-60x over the quiet floor, also near the threshold
        ln = int(rng.integers(30, 400))
        band[f:f + ln] *= np.float32(rng.uniform(8, 60))
        f += ln + int(rng.integers(200, 3000))
    g_met, g_segs = fv.VADMachine(fv.vad_machine_config(**over), SR, FFT).scan(band[None, :], ratio, frames_per_launch=step)
    o_met, o_segs = _oracle_scan(po, po.VadMachineConfig.default(**over), band, ratio)
    assert 1000 < o_met.sum() < n - 1000
    assert np.array_equal(g_met[0], o_met), int((g_met[0] != o_met).sum())
    assert [(a, b) for a, b, _, _ in g_segs[0]] == [(a, b) for a, b, _, _ in o_segs] and len(o_segs) > 10
