#!/usr/bin/env python
"""bench.py -- realtime factor (audio-sec/sec) of the full STFT -> NSNet2 -> ISTFT -> band energy -> VAD
hot path on B200, plus the STFT kernel's HBM GB/s and the NSNet2 roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W]               # this repo's CUDA path
    python bench.py --impl reference [...]                            # the reference's CPU algorithm (oracle port)

A "step" is one pass of the hot path over one batch: `--streams` (20) independent mono streams of
`--hours` (2) h each per GPU (BASELINE.json configs[2]; config 4 = the same 20 streams on each of
N GPUs, weak scaling, no collective on the data path).  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

# algorithmic work per unit.  Unit = one NSNet2 frame hop (10 ms of one channel).
FLOP_PER_FRAME_STEP = 2 * 2_681_000          # fc1 + 2 x (GRU W + GRU R) + fc2 + fc3 + fc4, MAC x 2 (as the reference executes it)
MAC_FC1, MAC_GRU_W, MAC_GRU_R, MAC_TAIL = 64_400, 480_000, 480_000, 240_000 + 360_000 + 96_600
MAC_FOLDED_W1 = 161 * 1200                   # fc1 folded into the first GRU's input projection (nn_tc.cu): one K = 161 layer
STEPS_PER_CHUNK, FRAMES_PER_CHUNK = 54, 50   # the reference evaluates 54 rows per 50 output frames
ISTFT_BYTES_PER_FRAME = 161 * 8 + 161 * 4 + 160 * 4  # spectrum row + f32 gain row read, one hop of y16 written


def stft_bytes_per_frame(D: int, with_spec: bool, fp16_rows: bool = False) -> int:
    # read 160*D new input samples, write 161 features (f32, or a 176-half row for the tensor-core backend)
    # (+ 161 complex bins when the spectrum is materialised)
    return 160 * D * 4 + (176 * 2 if fp16_rows else 161 * 4) + (161 * 8 if with_spec else 0)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1400.0, "fallback"  # B200_PROFILING.md fallback figures


def _traffic_file(kernel: str):
    """the newest committed `ncu --set full` capture (profiles/rNN_traffic.json, written by tools/ncu_traffic.py) that
    holds `kernel`"""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            with open(p) as f:
                d = json.load(f)
            if kernel in d.get("kernels", {}):
                return d
    return None


def ncu_traffic(kernel: str, streams: int, hours: float, sample_rate: int):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of one launch of `kernel`, from the committed
    `ncu --set full` capture of this same workload; None when no capture of this workload is on file."""
    d = _traffic_file(kernel)
    if not d:
        return None
    w = d.get("workload", {})
    if (w.get("streams"), w.get("hours"), w.get("sample_rate")) != (streams, hours, sample_rate):
        return None
    return d["kernels"][kernel].get("dram_bytes_per_launch")


def ncu_traffic_chunks(kernel: str):
    """chunks per stream the captured launch covered (the first, full-size span of a step)"""
    d = _traffic_file(kernel)
    return d["kernels"][kernel].get("chunks_per_stream") if d else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe).  The sampler is started
    before the warm-up steps (nvidia-smi needs a second or more to come up on an 8-GPU box, longer than a short timed
    region) and every line is stamped when it is read; stop(t0, t1) keeps the lines that fall inside the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t_begin: float = 0.0, t_end: float = float("inf")) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for ts, ln in self.lines if t_begin <= ts <= t_end + 0.03]
        window = "timed region"
        if not inside:  # sampling slower than the region: the warm-up steps are the same workload at the same load
            inside, window = [ln for _, ln in self.lines], "warm-up + timed region"
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples whose power is above the median idle-ish floor
        thr = 0.5 * max(pw)
        load = [s for s, p in zip(sm, pw) if p >= thr] or sm
        return {"sm_mhz": float(np.median(load)), "sm_mhz_min": float(min(load)), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "power_w_median": float(np.median([p for p in pw if p >= thr] or pw)), "samples": len(sm), "window": window,
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference pipeline, one OS thread per stream instance
# (reference simulator.zig:221-232), NN single-threaded per channel (onnxruntime.zig:58).
# ------------------------------------------------------------------------------------------------
def cpu_pipeline_throughput(fv, wd, sample_rate: int, n_threads: int, seconds_each: float, repeats: int = 1):
    from oracle import pyoracle as po
    po.lib()
    xs = [fv.synth.onboard_stream(9000 + i, seconds_each, 16000, 1)[0] for i in range(n_threads)]
    if sample_rate == 48000:
        xs = [fv.synth.expand_16k_to_48k(x) for x in xs]
    times = []
    nseg = 0
    for _ in range(repeats):
        pipes = [po.Pipeline(wd, sample_rate, 1) for _ in range(n_threads)]

        def work(i):
            pipes[i].push_samples(xs[i])  # ctypes releases the GIL for the call

        th = [threading.Thread(target=work, args=(i,)) for i in range(n_threads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        times.append(time.perf_counter() - t0)
        nseg = sum(len(p.segments(0)) for p in pipes)
        for p in pipes:
            p.close()
    audio = n_threads * seconds_each
    return audio, times, nseg


_JSON_FD = None
_ORIG_AFFINITY = None


def protect_stdout():
    """The contract is ONE JSON line on stdout; libraries (NCCL prints its version banner there) must not add to it:
    everything else that writes to fd 1 goes to stderr, the JSON line goes to the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(out: dict):
    line = (json.dumps(out) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def bind_to_gpu_numa(local: int):
    """Pin this rank to the CPUs next to its GPU (sysfs local_cpulist of the GPU's PCI function) before any pinned
    host memory is allocated, so first-touch places the staging buffers on the GPU's own NUMA node.  Returns the cpulist
    used, or None when it cannot be determined (the run continues unpinned)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local)).busId
        if isinstance(bus, bytes):
            bus = bus.decode()
        bus = bus.lower()
        if bus.count(":") == 2 and len(bus.split(":")[0]) == 8:
            bus = bus[4:]  # nvml prints an 8-digit domain, sysfs a 4-digit one
        with open(f"/sys/bus/pci/devices/{bus}/local_cpulist") as f:
            cpulist = f.read().strip()
        cpus = set()
        for part in cpulist.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        global _ORIG_AFFINITY
        _ORIG_AFFINITY = os.sched_getaffinity(0)
        cpus &= _ORIG_AFFINITY
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpulist
    except Exception:  # noqa: BLE001 -- best effort; an unpinned run is still a valid run
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    import fvad_b200
    fv = fvad_b200.load()
    wd = fv.weights.synthetic(seed=1234)
    cores = os.cpu_count() or 1
    sec_each = args.ref_seconds
    for _ in range(args.warmup):
        cpu_pipeline_throughput(fv, wd, args.sample_rate, cores, min(sec_each, 10.0))
    audio, times, _ = cpu_pipeline_throughput(fv, wd, args.sample_rate, cores, sec_each, repeats=args.steps)
    total = sum(times)
    value = audio * args.steps / total
    sample = f"{cores} streams x {sec_each:.0f} s mono {args.sample_rate} Hz per step (bounded sample of the workload)"
    out = {
        "impl": "reference", "metric": "realtime_factor_full_pipeline", "value": value, "unit": "audio-sec/sec",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "audio-sec/sec", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-sec/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(out)


def workload_config(args):
    return {"workload": f"full pipeline (STFT->NSNet2 GRU->ISTFT->band energy->VAD), {args.streams} streams x "
                        f"{args.hours:g} h per GPU, mono {args.sample_rate} Hz, synthetic NSNet2-baseline-shaped weights",
            "streams_per_gpu": args.streams, "hours": args.hours, "sample_rate": args.sample_rate,
            "nn_mode": args.nn_mode, "cache": "inputs (>= 9 GB per GPU) larger than L2; no flush needed",
            "parallelism": f"stream-sharded x{args.gpus}, no collective"}


def torch_cpu_nn_standin(wd, calls: int = 20):
    """the oracle's hand-written GEMV under-represents what onnxruntime/MLAS does on one core, so the NSNet2
    call of the reference (one [1,54,161] tensor per 0.5 s chunk, ONE intra-op thread: onnxruntime.zig:58,85-88) is also
    timed with torch's CPU kernels at one thread.  Returns ms per call and the audio-seconds per second one core sustains
    on the network alone."""
    import torch
    H = wd["gru1_R"].shape[1]
    prev = torch.get_num_threads()
    torch.set_num_threads(1)
    try:
        with torch.no_grad():
            t = lambda k: torch.from_numpy(np.ascontiguousarray(wd[k]))
            grus = []
            for l in (1, 2):
                g = torch.nn.GRU(H, H, batch_first=True)
                # ONNX gate order z,r,h -> torch r,z,n
                perm = torch.cat([torch.arange(H, 2 * H), torch.arange(0, H), torch.arange(2 * H, 3 * H)])
                g.weight_ih_l0.copy_(t(f"gru{l}_W")[perm]); g.weight_hh_l0.copy_(t(f"gru{l}_R")[perm])
                g.bias_ih_l0.copy_(t(f"gru{l}_Wb")[perm]); g.bias_hh_l0.copy_(t(f"gru{l}_Rb")[perm])
                grus.append(g)
            x = torch.randn(1, STEPS_PER_CHUNK, wd["fc1_w"].shape[1])

            def fwd(x):
                y = torch.nn.functional.linear(x, t("fc1_w"), t("fc1_b"))
                if wd.get("fc1_relu"):
                    y = torch.relu(y)
                y, _ = grus[0](y)
                y, _ = grus[1](y)
                y = torch.relu(torch.nn.functional.linear(y, t("fc2_w"), t("fc2_b")))
                y = torch.relu(torch.nn.functional.linear(y, t("fc3_w"), t("fc3_b")))
                return torch.sigmoid(torch.nn.functional.linear(y, t("fc4_w"), t("fc4_b")))
            for _ in range(3):
                fwd(x)
            t0 = time.perf_counter()
            for _ in range(calls):
                fwd(x)
            ms = 1e3 * (time.perf_counter() - t0) / calls
    finally:
        torch.set_num_threads(prev)
    return {"ms_per_call_54x161": ms, "threads": 1, "audio_sec_per_sec_per_core_nn_only": 0.5 / (ms * 1e-3),
            "what": "torch CPU kernels, one thread, same weights: the reference's onnx_instance.run() per 0.5 s channel-chunk"}


def run_48k_stereo(fv, wd, nn_mode, local: int, seeds, streams: int, hours: float, steps: int):
    """The reference-legal shape (VADPipeline.zig:55-58 accepts 48 kHz only; main.zig:135-142 runs 2 channels, i.e. two
    NSNet2 sequences per stream-chunk): device-resident and host-f32 end-to-end realtime factor, audio seconds counted
    per STREAM.  A short workload next to the headline one (synthetic; channel 1 = channel 0 mixed with a neighbour
    stream, so speech is common to both channels and the noise is not)."""
    import torch
    sr, C = 48000, 2
    x16, _ = fv.synth.onboard_batch_torch(seeds[:streams], hours * 3600.0, 16000, f"cuda:{local}")
    nxt = torch.cat([x16[:, 1:], x16[:, -1:]], dim=1)
    x = torch.empty((streams, C, x16.shape[1] * 3), dtype=torch.float32, device=x16.device)
    x[:, 0, 0::3] = x16
    x[:, 0, 1::3] = x16 + (nxt - x16) * (1.0 / 3.0)
    x[:, 0, 2::3] = x16 + (nxt - x16) * (2.0 / 3.0)
    x[:, 1] = 0.8 * x[:, 0] + 0.2 * torch.roll(x[:, 0], 1, dims=0)
    del x16, nxt
    n = x.shape[2]
    n_chunks = n // 24000
    audio_sec = streams * n_chunks * 24000 / sr
    gp = fv.BatchPipeline(streams, sr, C, wd, nn_mode=nn_mode, device=local)
    ptrs = [x[s, c].data_ptr() for s in range(streams) for c in range(C)]
    for _ in range(3):
        gp.push_pointers(ptrs, n, on_device=True)
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(steps):
        gp.push_pointers(ptrs, n, on_device=True)
        ms += gp.stats()["last_push_device_ms"]
    ms /= steps
    xh = torch.empty(x.shape, dtype=torch.float32, pin_memory=True)
    xh.copy_(x)
    torch.cuda.synchronize()
    hp = [xh[s, c].data_ptr() for s in range(streams) for c in range(C)]
    gp.push_pointers(hp, n, on_device=False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        gp.push_pointers(hp, n, on_device=False)
        segs = [gp.segments(s) for s in range(streams)]
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / steps
    gp.close()
    return {"workload": f"full pipeline, {streams} streams x {hours:g} h, 48000 Hz x {C} channels (the shape the reference accepts: "
                        f"{streams * C} NSNet2 sequences per 0.5 s)", "value": audio_sec / (ms * 1e-3), "unit": "stream-audio-sec/sec",
            "ms_per_step": ms, "steps": steps, "segments_found": int(sum(len(g) for g in segs)),
            "e2e": {"value": audio_sec / (e2e_ms * 1e-3), "unit": "stream-audio-sec/sec", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(streams * C * n * 4), "input": "pinned host f32 planar, double-buffered H2D"}}


def run_native(args):
    import torch
    import torch.distributed as dist
    import fvad_b200
    fv = fvad_b200.load()
    capi = fv.capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    host_cpus = bind_to_gpu_numa(local) if not os.environ.get("FVAD_NO_NUMA_BIND") else None
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sr = args.sample_rate
    D = sr // 16000
    S = args.streams
    seconds = args.hours * 3600.0
    chunk = 8000 * D
    wd = fv.weights.synthetic(seed=1234)
    nn_mode = {"f32": capi.NN_F32, "f16x3": capi.NN_TC_F16X3, "f16": capi.NN_TC_F16}[args.nn_mode]

    # ---- synthetic input, resident in HBM (and, for e2e, in pinned host memory) ----
    seeds = [1000 * (rank + 1) + s for s in range(S)]
    x16, labels = fv.synth.onboard_batch_torch(seeds, seconds, 16000, f"cuda:{local}")
    if D == 3:
        x = torch.empty((S, x16.shape[1] * 3), dtype=torch.float32, device=x16.device)
        nxt = torch.cat([x16[:, 1:], x16[:, -1:]], dim=1)
        x[:, 0::3] = x16
        x[:, 1::3] = x16 + (nxt - x16) * (1.0 / 3.0)
        x[:, 2::3] = x16 + (nxt - x16) * (2.0 / 3.0)
        del nxt, x16
    else:
        x = x16
    n = x.shape[1]
    n_chunks = n // chunk
    audio_sec_per_step = S * n_chunks * chunk / sr  # only whole chunks are processed (VADPipeline.zig:150)
    dev_ptrs = [x[s].data_ptr() for s in range(S)]

    gp = fv.BatchPipeline(S, sr, 1, wd, nn_mode=nn_mode, device=local)
    clocks = ClockSampler(local)

    # ---- device-resident arm ----
    clocks.start()
    for _ in range(args.warmup):
        gp.push_pointers(dev_ptrs, n, on_device=True)
    st0 = gp.stats()
    barrier()
    t0 = time.perf_counter()
    dev_ms = 0.0
    stage = np.zeros(5)
    layer_ms, layer_n = np.zeros(9), np.zeros(9)
    for _ in range(args.steps):
        gp.push_pointers(dev_ptrs, n, on_device=True)
        st = gp.stats()
        dev_ms += st["last_push_device_ms"]
        stage += [st["last_stft_ms"], st["last_nn_ms"], st["last_istft_ms"], st["last_vadfft_ms"], st["last_scan_ms"]]
        layer_ms += st["last_nn_layer_ms"]
        layer_n += st["last_nn_layer_launches"]
    barrier()
    t1 = time.perf_counter()
    wall_ms = 1e3 * (t1 - t0)
    clk = clocks.stop(t0, t1)
    st1 = gp.stats()
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    n_segments = sum(len(gp.segments(s)) for s in range(S))

    # ---- end-to-end arm: host (pinned) PCM in, segments out, through the C ABI ----
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    xh = torch.empty((S, n), dtype=torch.float32, pin_memory=True)
    xh.copy_(x)
    torch.cuda.synchronize()
    host_ptrs = [xh[s].data_ptr() for s in range(S)]
    gp2 = fv.BatchPipeline(S, sr, 1, wd, nn_mode=nn_mode, device=local)
    gp2.push_pointers(host_ptrs, n, on_device=False)  # warm-up (allocations)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gp2.push_pointers(host_ptrs, n, on_device=False)
        segs = [gp2.segments(s) for s in range(S)]
    barrier()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    d2h = sum(len(sg) for sg in segs) * 24 + S * 128
    del xh, host_ptrs  # the 16-bit arm reuses the pinned block (8 ranks x 9 GB of pinned memory is enough)
    # same, 16-bit PCM host buffers (what a WAV holds; fvad_batch_push_i16 converts on the device)
    xq = torch.empty((S, n), dtype=torch.int16, pin_memory=True)
    xq.copy_((x * 32768.0).round_().clamp_(-32768, 32767).to(torch.int16))
    torch.cuda.synchronize()
    q_ptrs = [xq[s].data_ptr() for s in range(S)]
    gp2.push_pointers(q_ptrs, n, on_device=False, i16=True)  # warm-up (allocations)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        gp2.push_pointers(q_ptrs, n, on_device=False, i16=True)
        segs16 = [gp2.segments(s) for s in range(S)]
    barrier()
    e2e16_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    gp2.close()

    # ---- the link the e2e arm is bound by: plain pinned-host -> device copies of the same buffer, all ranks at once ----
    link_bytes = min(xq.numel() * 2, 2 << 30)
    src = xq.view(-1)[:link_bytes // 2]
    dst = torch.empty(link_bytes // 2, dtype=torch.int16, device=f"cuda:{local}")
    dst.copy_(src, non_blocking=True)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    ev1.record()
    barrier()
    link_gbs = 3 * link_bytes / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    del dst, src, xq

    # max over ranks
    tt = torch.tensor([dev_ms, wall_ms, e2e_ms, e2e16_ms, -link_gbs], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms, e2e16_ms, link_gbs = [float(v) for v in tt.tolist()]
    link_gbs = -link_gbs  # the slowest rank's link

    if rank == 0:
        hbm_peak, tf_peak, which = measured_peaks()
        ms_step = dev_ms / args.steps
        value = world * audio_sec_per_step / (ms_step * 1e-3)
        frames = S * n_chunks * FRAMES_PER_CHUNK
        nn_flop_ref = S * n_chunks * STEPS_PER_CHUNK * FLOP_PER_FRAME_STEP  # what the reference's ORT run executes
        # what this backend executes: fc1 folded into W1 when the graph has no activation between them (tensor-core
        # backend), fc2..fc4 only on the 50 rows whose output the reference keeps (NSNet2.zig:188-190)
        folded = args.nn_mode != "f32" and not wd.get("fc1_relu") and os.environ.get("FVAD_NN_FOLD_FC1", "1") != "0"
        mac_rec = (MAC_FOLDED_W1 if folded else MAC_FC1 + MAC_GRU_W) + MAC_GRU_W + 2 * MAC_GRU_R
        nn_flop = 2.0 * S * n_chunks * (STEPS_PER_CHUNK * mac_rec + FRAMES_PER_CHUNK * MAC_TAIL)
        gru_kernel = "tc_gru_kernel" if os.environ.get("FVAD_GRU_PAIR", "1") == "0" else "tc_gru2_kernel"
        # the first GRU layer with its (folded) input projection fused in (nn_tc_gru3.cu) is then the dominant kernel
        fused = folded and gru_kernel == "tc_gru2_kernel" and os.environ.get("FVAD_GRU_FUSE", "1") != "0"
        nn_ms = stage[1] / args.steps
        stft_ms = stage[0] / args.steps
        istft_ms = stage[2] / args.steps  # the stage also holds two small carry copies (kilobytes)
        ach_tf = nn_flop / (nn_ms * 1e-3) / 1e12 if nn_ms > 0 else 0.0
        sb = frames * stft_bytes_per_frame(D, True, args.nn_mode != "f32")
        ach_gbs = sb / (stft_ms * 1e-3) / 1e9 if stft_ms > 0 else 0.0
        # dominant kernel: one GRU layer launch (tc_gru_kernel; two layers, identical shapes).  Algorithmic flops of a
        # launch = its share of S*n_chunks sequences x 54 steps x 2*480000 (R.h of the layer).
        if fused:  # one launch = R.h of layer 1 + the folded projection (161 x 1200) over all 54 steps of its tiles
            gru_kernel = "tc_gru3_kernel"
            gru_ms, gru_n = layer_ms[3], layer_n[3]
            gru_flop_total = 2.0 * args.steps * S * n_chunks * STEPS_PER_CHUNK * (MAC_GRU_R + MAC_FOLDED_W1)
        else:
            gru_ms, gru_n = layer_ms[3] + layer_ms[5], layer_n[3] + layer_n[5]
            gru_flop_total = 2.0 * args.steps * S * n_chunks * STEPS_PER_CHUNK * 2 * MAC_GRU_R
        gru_tf = gru_flop_total / (gru_ms * 1e-3) / 1e12 if gru_ms > 0 else 0.0
        names = fv.capi.NN_LAYERS
        kc = ncu_traffic_chunks("stft_kernel_d1")  # algorithmic bytes of the launch the ncu traffic figure belongs to
        stft_traffic_alg = (S * kc * FRAMES_PER_CHUNK * stft_bytes_per_frame(D, True, args.nn_mode != "f32")
                            if kc and ncu_traffic("stft_kernel_d1", S, args.hours, args.sample_rate) else None)
        out = {
            "metric": "realtime_factor_full_pipeline", "value": value, "unit": "audio-sec/sec", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "wall_ms_per_step": wall_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"f32": "f32", "f16x3": "f16x3-split (fp32 accumulate)", "f16": "f16 (fp32 accumulate)"}[args.nn_mode],
            "data": "synthetic", "config": workload_config(args),
            "stage_ms_per_step": {k: float(v) / args.steps for k, v in zip(["stft", "nsnet2", "istft", "vadfft", "vadscan"], stage)},
            "nn_stage_roofline": {"kernel": "nsnet2 (GEMMs + 2x54-step GRU)", "bound": "tensor", "achieved": ach_tf, "peak": tf_peak,
                         "unit": "TFLOP/s", "frac": ach_tf / tf_peak, "traffic": None,  # a stage of eight kernels; per-kernel traffic below
                         "flop_counted": "executed: " + ("fc1 folded into the first GRU input projection (K = 161)"
                                                         + (", itself fused into the first GRU kernel, " if fused else ", ") if folded else "")
                                         + "fc2-fc4 on the 50 kept rows of 54",
                         "flop_executed": nn_flop, "flop_as_executed_by_reference": float(nn_flop_ref),
                         "achieved_on_reference_count": nn_flop_ref / (nn_ms * 1e-3) / 1e12 if nn_ms > 0 else 0.0,
                         "peak_source": f"{which} bf16_tflops_sustained"},
            "roofline": {"kernel": gru_kernel + (" (dominant: first GRU layer + its input projection, %d launches per step)" % int(gru_n / max(1, args.steps))
                                                 if fused else " (dominant: 2 layers x %d launches per step)" % int(gru_n / max(1, 2 * args.steps))),
                                "bound": "tensor", "achieved": gru_tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": gru_tf / tf_peak,
                                "avg_launch_ms": gru_ms / max(gru_n, 1), "flop_per_launch": gru_flop_total / max(gru_n, 1),
                                "traffic": ncu_traffic(gru_kernel, S, args.hours, args.sample_rate),
                                "peak_source": f"{which} bf16_tflops_sustained"},
            "nn_layer_ms_per_step": {k: float(v) / args.steps for k, v in zip(names, layer_ms)},
            "stft_roofline": {"kernel": "stft_kernel", "bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s",
                              "frac": ach_gbs / hbm_peak, "traffic": ncu_traffic("stft_kernel_d1", S, args.hours, args.sample_rate),
                              "traffic_launch_algorithmic_bytes": stft_traffic_alg, "bytes_per_frame": stft_bytes_per_frame(D, True, args.nn_mode != "f32"),
                              "peak_source": f"{which} hbm_gbs"},
            "istft_roofline": {"kernel": "istft_stream_kernel", "bound": "hbm",
                               "achieved": frames * ISTFT_BYTES_PER_FRAME / (istft_ms * 1e-3) / 1e9 if istft_ms > 0 else 0.0,
                               "peak": hbm_peak, "unit": "GB/s",
                               "frac": (frames * ISTFT_BYTES_PER_FRAME / (istft_ms * 1e-3) / 1e9 / hbm_peak) if istft_ms > 0 else 0.0,
                               "traffic": ncu_traffic("istft_stream_kernel", S, args.hours, args.sample_rate),
                               "traffic_launch_algorithmic_bytes": (S * kc * FRAMES_PER_CHUNK * ISTFT_BYTES_PER_FRAME
                                                                    if kc and ncu_traffic("istft_stream_kernel", S, args.hours, args.sample_rate) else None),
                               "bytes_per_frame": ISTFT_BYTES_PER_FRAME, "peak_source": f"{which} hbm_gbs"},
            "e2e": {"value": world * audio_sec_per_step / (e2e_ms * 1e-3), "unit": "audio-sec/sec",
                    "h2d_bytes_per_step": int(S * n * 4), "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms,
                    "steps": e2e_steps, "input": "pinned host f32 planar (the reference's pushSamples type), double-buffered H2D",
                    "h2d_link_gbs": link_gbs, "h2d_achieved_gbs": S * n * 4 / (e2e_ms * 1e-3) / 1e9,
                    "h2d_frac_of_link": (S * n * 4 / (e2e_ms * 1e-3) / 1e9) / link_gbs if link_gbs > 0 else None,
                    "link_note": "h2d_link_gbs = plain pinned-host -> device copies of 2 GiB on every rank at once "
                                 "(slowest rank); the f32 arm is bound by it"},
            "e2e_i16": {"value": world * audio_sec_per_step / (e2e16_ms * 1e-3), "unit": "audio-sec/sec",
                        "h2d_bytes_per_step": int(S * n * 2), "ms_per_step": e2e16_ms, "steps": e2e_steps,
                        "input": "pinned host int16 PCM, converted on the device"},
            "gpu_launches": int(launches), "segments_found": int(n_segments), "clocks": clk,
            "host_affinity": host_cpus,
        }
        if world == 1 and not args.no_48k:
            del x
            torch.cuda.empty_cache()
            out["ref_legal_48k_stereo"] = run_48k_stereo(fv, wd, nn_mode, local, seeds, S, min(args.hours, 0.25), 2)
        # CPU baseline on a bounded sample (N=1 only)
        if world == 1 and not args.no_cpu_baseline:
            if _ORIG_AFFINITY:
                os.sched_setaffinity(0, _ORIG_AFFINITY)  # the CPU arm uses every host core
            cores = os.cpu_count() or 1
            audio, times, _ = cpu_pipeline_throughput(fv, wd, sr, cores, args.cpu_seconds)
            out["cpu_baseline"] = {"value": audio / times[0], "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                                   "sample": f"{cores} streams x {args.cpu_seconds:.0f} s, one thread per stream (oracle port of the "
                                             f"reference pipeline; reference binary unavailable: no zig/onnxruntime on this box)",
                                   "library_grade_nn": torch_cpu_nn_standin(wd)}
        emit(out)
    gp.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--streams", type=int, default=20)
    ap.add_argument("--hours", type=float, default=2.0)
    ap.add_argument("--sample-rate", type=int, default=16000, choices=[16000, 48000])
    ap.add_argument("--nn-mode", default="f16", choices=["f32", "f16x3", "f16"])
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-seconds", type=float, default=200.0)
    ap.add_argument("--ref-seconds", type=float, default=120.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-48k", action="store_true", help="skip the extra 48 kHz x 2 channel line (N = 1 only)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "native":
        print("bench.py: warning: W < 3 warm-up steps breaks the timing rules", file=sys.stderr)
    protect_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
