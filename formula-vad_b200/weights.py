"""NSNet2 weight container: seeded synthetic generator and the on-disk blob format.

The reference loads `data/nsnet2-20ms-baseline.onnx` (src/NSNet2.zig:53-58); that file and
onnxruntime are absent from the reference snapshot (.MISSING_LARGE_BLOBS), so tests and the
benchmark run on seeded synthetic weights of the public NSNet2-baseline shapes:
fc1 161->400, GRU 400->400 x2 (ONNX layout, gate order z,r,h, linear_before_reset=1),
fc2 400->600 ReLU, fc3 600->600 ReLU, fc4 600->161 sigmoid.

Blob layout (little endian), read by csrc/nsnet2_weights.cpp:
    char[8]  "FVADNS2\\0"
    u32      version (1)
    u32      fc1_relu
    u32      n_bins, hidden, fc2, fc3
    f32[]    fc1_w[H][B] fc1_b[H]
             gru1_W[3H][H] gru1_R[3H][H] gru1_Wb[3H] gru1_Rb[3H]
             gru2_W ... (same)
             fc2_w[F2][H] fc2_b[F2] fc3_w[F3][F2] fc3_b[F3] fc4_w[B][F3] fc4_b[B]
"""
from __future__ import annotations

import struct

import numpy as np

MAGIC = b"FVADNS2\0"
NAMES = ("fc1_w", "fc1_b", "gru1_W", "gru1_R", "gru1_Wb", "gru1_Rb", "gru2_W", "gru2_R",
         "gru2_Wb", "gru2_Rb", "fc2_w", "fc2_b", "fc3_w", "fc3_b", "fc4_w", "fc4_b")


def shapes(n_bins=161, hidden=400, fc2=600, fc3=600):
    H = hidden
    return {"fc1_w": (H, n_bins), "fc1_b": (H,),
            "gru1_W": (3 * H, H), "gru1_R": (3 * H, H), "gru1_Wb": (3 * H,), "gru1_Rb": (3 * H,),
            "gru2_W": (3 * H, H), "gru2_R": (3 * H, H), "gru2_Wb": (3 * H,), "gru2_Rb": (3 * H,),
            "fc2_w": (fc2, H), "fc2_b": (fc2,), "fc3_w": (fc3, fc2), "fc3_b": (fc3,),
            "fc4_w": (n_bins, fc3), "fc4_b": (n_bins,)}


def synthetic(seed: int = 1234, n_bins=161, hidden=400, fc2=600, fc3=600, fc1_relu=0) -> dict:
    """U(-1/sqrt(fan_in), 1/sqrt(fan_in)) like torch's default init; seed recorded by callers."""
    rng = np.random.default_rng(seed)
    out = {"fc1_relu": int(fc1_relu)}
    for name, shp in shapes(n_bins, hidden, fc2, fc3).items():
        if name.startswith("gru"):
            fan = hidden
        elif name.endswith("_w"):
            fan = shp[1]
        else:  # bias: fan_in of its layer
            fan = {"fc1_b": n_bins, "fc2_b": hidden, "fc3_b": fc2, "fc4_b": fc3}[name]
        k = 1.0 / np.sqrt(fan)
        out[name] = rng.uniform(-k, k, size=shp).astype(np.float32)
    return out


def save_blob(path: str, wd: dict) -> None:
    H, B = wd["fc1_w"].shape
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<6I", 1, int(wd.get("fc1_relu", 0)), B, H, wd["fc2_w"].shape[0], wd["fc3_w"].shape[0]))
        for n in NAMES:
            f.write(np.ascontiguousarray(wd[n], np.float32).tobytes())


def load_blob(path: str) -> dict:
    with open(path, "rb") as f:
        if f.read(8) != MAGIC:
            raise ValueError("not an FVADNS2 weight blob")
        ver, relu, B, H, F2, F3 = struct.unpack("<6I", f.read(24))
        if ver != 1:
            raise ValueError(f"unsupported blob version {ver}")
        out = {"fc1_relu": relu}
        for n, shp in shapes(B, H, F2, F3).items():
            cnt = int(np.prod(shp))
            out[n] = np.frombuffer(f.read(4 * cnt), np.float32).reshape(shp).copy()
    return out
