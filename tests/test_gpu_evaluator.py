"""Evaluator on the device (csrc/eval.cu) through the C ABI: the reference's own known answers
(src/Evaluator/statistics.zig:286-360), bit-for-bit parity with the oracle on random segment lists, and the
batched form scoring a threshold/hysteresis sweep (BASELINE config 5 in miniature)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FIELDS = ("total_positives_sec", "true_positives_sec", "false_positives_sec", "false_negatives_sec", "true_positive_rate",
          "precision", "false_negative_rate", "false_discovery_rate", "f_score", "fm_index")
ORC = ("p_sec", "tp_sec", "fp_sec", "fn_sec", "tpr", "ppv", "fnr", "fdr", "f_score", "fm")


def _oracle_stats(po, vad, ref, cfg):
    L = po.lib()
    v = (po.EvalSeg * max(len(vad), 1))(*[po.EvalSeg(a, b) for a, b in vad])
    r = (po.EvalSeg * max(len(ref), 1))(*[po.EvalSeg(a, b) for a, b in ref])
    oc = po.EvalCfg(cfg.extrude_start, cfg.extrude_end, cfg.fill_gaps, cfg.ignore_shorter_than_sec, cfg.f_score_beta)
    st = po.EvalStats()
    L.orc_evaluate(v, len(vad), r, len(ref), C.byref(oc), C.byref(st))
    return {f: getattr(st, o) for f, o in zip(FIELDS, ORC)}


def _same(a, b):
    for f in FIELDS:
        x, y = np.float32(a[f]), np.float32(b[f])
        assert (np.isnan(x) and np.isnan(y)) or x == y, (f, x, y)


def test_reference_known_answers(fv):
    """calcFalsePositiveSec #1 / #2: one VAD segment against refs (2,3), (4,5), extrusion 2/2, gap fill 2 -> FP 0 s / 3 s."""
    cfg = fv.eval_config(extrude_start=2, extrude_end=2, fill_gaps=2)
    refs = [(2, 3), (4, 5)]
    st = fv.Evaluator.init_and_run([(1, 6)], refs).stats(cfg)
    assert abs(st["false_positives_sec"] - 0.0) < 1e-3 and abs(st["true_positives_sec"] - 5.0) < 1e-3
    st = fv.Evaluator.init_and_run([(1, 10)], refs).stats(cfg)
    assert abs(st["false_positives_sec"] - 3.0) < 1e-3 and abs(st["true_positives_sec"] - 6.0) < 1e-3


def test_hand_checked_rates_and_empty_lists(fv):
    st = fv.Evaluator.init_and_run([(50, 55), (10, 20)], [(80, 90), (12, 18)]).stats()  # unsorted on purpose
    assert abs(st["true_positives_sec"] - 6) < 1e-3 and abs(st["false_positives_sec"] - 9) < 1e-3
    assert abs(st["false_negatives_sec"] - 10) < 1e-3 and abs(st["total_positives_sec"] - 16) < 1e-3
    assert abs(st["true_positive_rate"] - 6 / 16) < 1e-4 and abs(st["precision"] - 6 / 15) < 1e-4
    st = fv.Evaluator.init_and_run([], [(1, 2)]).stats()
    assert st["false_negatives_sec"] == 1.0 and st["true_positive_rate"] == 0.0 and np.isnan(st["precision"])
    st = fv.Evaluator.init_and_run([], []).stats()
    assert np.isnan(st["true_positive_rate"])  # 0/0 like the reference's f32 division


def test_random_lists_match_oracle_bitwise(fv, po):
    rng = np.random.default_rng(5)
    for case in range(40):
        def segs(n, span):
            starts = np.sort(rng.uniform(0, span, n)).astype(np.float32)
            return [(float(s), float(np.float32(s + rng.uniform(0.2, 15.0)))) for s in starts]  # overlapping allowed
        vad, ref = segs(int(rng.integers(0, 30)), 600.0), segs(int(rng.integers(0, 30)), 600.0)
        cfg = fv.eval_config(ignore_shorter_than_sec=float(rng.choice([0, 1.0])), extrude_start=float(rng.choice([0, 2.0])),
                             extrude_end=float(rng.choice([0, 2.0, 5.0])), fill_gaps=float(rng.choice([0, 2.0, 4.0])))
        perm = rng.permutation(len(vad))
        got = fv.Evaluator.init_and_run([vad[i] for i in perm], ref).stats(cfg)
        _same(got, _oracle_stats(po, vad, ref, cfg))


def test_batch_sweep_scores_every_stream_and_machine(fv, po, wd):
    """Config 5 in miniature: alt machines sweep the threshold factor; every (stream, machine) list is scored on
    the device against the stream's labels and equals the oracle's Evaluator fed the same segments."""
    sr, S = 16000, 3
    factors = [2.0, 4.0, 6.0, 8.0, 12.0, 16.0, 24.0]
    xs, labels = zip(*[fv.synth.onboard_stream(300 + s, 60.0, sr, 1) for s in range(S)])
    gp = fv.BatchPipeline(S, sr, 1, wd, alt_vad_machine_configs=[fv.vad_machine_config(speech_threshold_factor=f) for f in factors])
    gp.push(np.concatenate(xs, axis=0))
    refs = [[(float(np.float32(a)), float(np.float32(b))) for a, b in lab] for lab in labels]
    cfg = fv.eval_config(extrude_start=2.0, extrude_end=2.0, fill_gaps=2.0)
    table = gp.evaluate(refs, cfg)
    assert len(table) == S and len(table[0]) == 1 + len(factors)
    n_lists = 0
    for s in range(S):
        for m in range(1 + len(factors)):
            segs = gp.segments(s, m)
            vad = [(float(np.float32(f) / np.float32(sr)), float(np.float32(t) / np.float32(sr))) for f, t, _, _ in segs]
            _same(table[s][m], _oracle_stats(po, vad, refs[s], cfg))
            n_lists += len(segs) > 0
    assert n_lists >= S  # the sweep does produce segments to score
    gp.close()
