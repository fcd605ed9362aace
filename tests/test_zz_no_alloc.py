"""fvad_vad_machine_scan_noalloc (no counterpart in the reference): the VAD
machine operator with its scratch in a static __device__ array, called through the C ABI with caller-owned device
buffers, must give exactly what fvad_vad_machine_scan gives -- on two calls in a row (the scratch is shared and keeps
the previous call's contents, so the second call, with other machines and another series, also proves the re-initialisation)
-- and must refuse sizes beyond its stated limits.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SR, FFT = 48000, 1024


def _series(n, seed):
    rng = np.random.default_rng(seed)
    band = (0.004 * np.exp(0.5 * rng.standard_normal(n))).astype(np.float32)
    ratio = np.clip(0.75 + 0.2 * rng.standard_normal(n), 0.0, 1.0).astype(np.float32)
    f = 300
    while f < n - 400:
        ln = int(rng.integers(25, 280))
        band[f:f + ln] *= np.float32(rng.uniform(30, 400))
        f += ln + int(rng.integers(40, 900))
    return band, ratio


def _noalloc_scan(fv, cfgs, band, ratio, max_segments):
    import torch
    L = fv.capi.lib()
    M, n = band.shape
    dev = torch.device("cuda", 0)
    d_band = torch.from_numpy(band).to(dev).contiguous()
    d_ratio = torch.from_numpy(ratio).to(dev).contiguous()
    d_met = torch.full((M, n), 255, dtype=torch.uint8, device=dev)
    d_seg = torch.zeros((M * max_segments * C.sizeof(fv.capi.Segment),), dtype=torch.uint8, device=dev)
    d_cnt = torch.full((M,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    arr = (fv.capi.VadMachineConfig * M)(*cfgs)
    rc = L.fvad_vad_machine_scan_noalloc(arr, M, SR, FFT, d_band.data_ptr(), d_ratio.data_ptr(), n, d_met.data_ptr(),
                                         d_seg.data_ptr(), max_segments, d_cnt.data_ptr(), None)
    fv.capi.check(rc)
    met = d_met.cpu().numpy()
    cnt = d_cnt.cpu().numpy()
    raw = d_seg.cpu().numpy().tobytes()
    segs = (fv.capi.Segment * (M * max_segments)).from_buffer_copy(raw)
    out = []
    for m in range(M):
        k = min(int(cnt[m]), max_segments)
        out.append([(segs[m * max_segments + i].sample_from, segs[m * max_segments + i].sample_to,
                     segs[m * max_segments + i].avg_channel_vol_ratio, segs[m * max_segments + i].vad_met_sec) for i in range(k)])
    return met, out, cnt


def test_noalloc_scan_equals_the_allocating_operator_twice_in_a_row(fv):
    calls = [
        # call 1: three machines (default, quick-to-open, no pre-filled long-term ring with a short window: generic kernel)
        ([dict(), dict(speech_threshold_factor=4.0, max_speech_gap_sec=0.5, min_vad_duration_sec=0.2),
          dict(has_initial_long_term_avg=0, long_term_speech_avg_sec=9.0)], 20000, 11),
        # call 2: other machines, another series, fewer frames: nothing of call 1 may survive in the scratch
        ([dict(speech_threshold_factor=20.0), dict(short_term_speech_avg_sec=0.05, long_term_speech_avg_sec=20.0)], 12345, 12),
    ]
    max_segments = 64
    for overs, n, seed in calls:
        cfgs = [fv.vad_machine_config(**o) for o in overs]
        rows = [_series(n, seed + i) for i in range(len(cfgs))]
        band = np.stack([b for b, _ in rows])
        ratio = rows[0][1]
        ref_met, ref_segs = fv.VADMachine(cfgs, SR, FFT, max_segments=max_segments).scan(band, ratio)
        met, segs, cnt = _noalloc_scan(fv, cfgs, band, ratio, max_segments)
        assert 0 < ref_met.sum() < ref_met.size and sum(len(s) for s in ref_segs) > 0  # the series exercises the machines
        assert np.array_equal(met, ref_met)          # identical decisions (the bar of tests/test_gpu_vad_machine.py)
        assert segs == ref_segs                      # identical segments, bit for bit
        assert [int(c) for c in cnt] == [len(s) for s in ref_segs]


def test_noalloc_scan_refuses_sizes_beyond_the_static_scratch(fv):
    import torch
    L = fv.capi.lib()
    buf = torch.zeros(1 << 16, dtype=torch.float32, device="cuda:0")
    cnt = torch.zeros(64, dtype=torch.int64, device="cuda:0")
    too_many = (fv.capi.VadMachineConfig * 17)(*[fv.vad_machine_config() for _ in range(17)])
    rc = L.fvad_vad_machine_scan_noalloc(too_many, 17, SR, FFT, buf.data_ptr(), buf.data_ptr(), 16, None, buf.data_ptr(), 4,
                                         cnt.data_ptr(), None)
    assert rc != 0 and b"FVAD_NOALLOC_MAX_MACHINES" in L.fvad_last_error_string()
    long_ring = (fv.capi.VadMachineConfig * 1)(fv.vad_machine_config(long_term_speech_avg_sec=600.0))  # 28 125 frames
    rc = L.fvad_vad_machine_scan_noalloc(long_ring, 1, SR, FFT, buf.data_ptr(), buf.data_ptr(), 16, None, buf.data_ptr(), 4,
                                         cnt.data_ptr(), None)
    assert rc != 0 and b"FVAD_NOALLOC_MAX_RING" in L.fvad_last_error_string()
