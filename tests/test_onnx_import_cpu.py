"""tools/onnx_to_blob.py: the self-contained ONNX reader must recover the NSNet2 weights from a graph file
(MatMul+Add and Gemm flavours, raw_data and float_data initializers), read activation placement and GRU
attributes from the node list, and refuse graphs the kernels do not implement.  The shipped
nsnet2-20ms-baseline.onnx is absent from the reference snapshot (.MISSING_LARGE_BLOBS), so the fixture
is written by the same wire-format code from seeded weights."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("onnx_to_blob", os.path.join(ROOT, "tools", "onnx_to_blob.py"))
o2b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(o2b)


@pytest.mark.parametrize("gemm,raw,relu", [(False, True, 0), (True, True, 1), (False, False, 0)])
def test_round_trip(fv, tmp_path, gemm, raw, relu):
    wd = fv.weights.synthetic(seed=77, fc1_relu=relu)
    path = str(tmp_path / "m.onnx")
    o2b.write_nsnet2_onnx(path, wd, gemm=gemm, raw=raw)
    blob = str(tmp_path / "m.fvadns2")
    got = o2b.convert(path, blob)
    assert got["fc1_relu"] == relu
    for k in fv.weights.NAMES:
        assert np.array_equal(got[k], wd[k]), k
    back = fv.weights.load_blob(blob)
    for k in fv.weights.NAMES:
        assert np.array_equal(back[k], wd[k]), k
    assert back["fc1_relu"] == relu


def test_torch_style_export_with_h0_chain(fv, tmp_path):
    """What torch.onnx.export writes for a batch_first model with h0 = None: initial_h comes from a
    Shape -> Gather -> Concat -> ConstantOfShape -> Slice chain, Squeeze takes `axes` as an input,
    Transpose(1,0,2) brackets the recurrent block.  The importer must accept it bit-exactly."""
    wd = fv.weights.synthetic(seed=91)
    path = str(tmp_path / "t.onnx")
    o2b.write_nsnet2_onnx(path, wd, torch_style=True)
    nodes, inits, gin, gout = o2b.parse_model(open(path, "rb").read())
    ops = [n["op"] for n in nodes]
    assert ops.count("ConstantOfShape") == 1 and ops.count("Slice") == 2 and ops.count("Transpose") == 2
    gru = [n for n in nodes if n["op"] == "GRU"]
    assert all(len(g["input"]) == 6 and g["input"][5] for g in gru)  # initial_h is bound, to a computed value
    got = o2b.extract_weights(nodes, inits, gin, gout)
    for k in fv.weights.NAMES:
        assert np.array_equal(got[k], wd[k]), k


def test_refuses_nonzero_h0_and_misplaced_frame_axis(fv, tmp_path):
    wd = fv.weights.synthetic(seed=92)
    path = str(tmp_path / "bad.onnx")
    # h0 filled with 0.25: not the reference's semantics (no state tensor is bound, NSNet2.zig:220)
    o2b.write_nsnet2_onnx(path, wd, torch_style=True, h0_value=0.25)
    with pytest.raises(ValueError, match="initial_h"):
        o2b.convert(path)
    # no Transpose before the GRUs: X = [1, 54, 400] would make batch the recurrence axis
    o2b.write_nsnet2_onnx(path, wd, drop_transposes=True)
    with pytest.raises(ValueError, match="sequence axis"):
        o2b.convert(path)
    # a Transpose that swaps frames and features
    o2b.write_nsnet2_onnx(path, wd)
    nodes, inits, gin, gout = o2b.parse_model(open(path, "rb").read())
    tr = next(n for n in nodes if n["op"] == "Transpose")
    tr["attr"]["perm"] = [2, 0, 1]
    with pytest.raises(ValueError, match="frames"):
        o2b.extract_weights(nodes, inits, gin, gout)
    # initial_h as a non-zero initializer
    o2b.write_nsnet2_onnx(path, wd)
    nodes, inits, gin, gout = o2b.parse_model(open(path, "rb").read())
    gru = next(n for n in nodes if n["op"] == "GRU")
    gru["input"] += ["", "h0"]
    inits["h0"] = np.full((1, 1, 400), 0.5, np.float32)
    with pytest.raises(ValueError, match="initial_h"):
        o2b.extract_weights(nodes, inits, gin, gout)
    inits["h0"] = np.zeros((1, 1, 400), np.float32)  # an explicit zero state is the reference's semantics
    assert o2b.extract_weights(nodes, inits, gin, gout)["fc1_relu"] == 0


def test_rejects_what_the_kernels_do_not_implement(fv, tmp_path):
    wd = fv.weights.synthetic(seed=3)
    path = str(tmp_path / "m.onnx")
    o2b.write_nsnet2_onnx(path, wd)
    data = open(path, "rb").read()
    nodes, inits, gin, gout = o2b.parse_model(data)
    assert [n["op"] for n in nodes].count("GRU") == 2 and gin == ["input"] and gout == ["output"]
    gru = next(n for n in nodes if n["op"] == "GRU")
    gru["attr"]["linear_before_reset"] = 0
    with pytest.raises(ValueError, match="linear_before_reset"):
        o2b.extract_weights(nodes, inits, gin, gout)
    gru["attr"]["linear_before_reset"] = 1
    sig = next(n for n in nodes if n["op"] == "Sigmoid")
    sig["op"] = "Tanh"
    with pytest.raises(ValueError, match="unsupported op"):
        o2b.extract_weights(nodes, inits, gin, gout)
    with pytest.raises(ValueError):
        o2b.parse_model(b"\x0a\x03abc")  # no graph
