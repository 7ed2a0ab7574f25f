"""Checkpoint loader and SmoothQuant calibration (host side, CPU): b200stage/loader.py.

The reference loads real checkpoints by path (`from_pretrained(model_path)`, Code/Quantised Models/
models_quant_updated.py:25-40).  No checkpoint exists offline, so the tests write one: the seeded synthetic
weights are loaded into the HF module the reference would instantiate, saved with `save_pretrained`, and read
back through the loader, one layer range at a time.
"""
import os
import sys

import numpy as np
import pytest
import torch

import b200stage
from b200stage import loader, weights
from oracle import w8a8 as O

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden  # noqa: E402


def _same(a: dict, b: dict):
    assert set(a) == set(b), (sorted(a), sorted(b))
    for k in a:
        assert a[k].dtype == b[k].dtype and torch.equal(a[k], b[k]), k


@pytest.mark.parametrize("model,safe", [("tiny-gqa-test", True), ("tiny-llama-test", False), ("tiny-gpt2-test", True),
                                        ("tiny-phi-test", True), ("tiny-neox-test", True)])
def test_layer_range_from_checkpoint_equals_direct_quantisation(tmp_path, model, safe):
    cfg0 = b200stage.get_config(model)
    seed = 3
    m = make_golden.hf_for(cfg0, seed)
    m.save_pretrained(tmp_path, safe_serialization=safe)
    cfg = b200stage.get_config(str(tmp_path))
    assert (cfg.n_layers, cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn, cfg.vocab) == \
           (cfg0.n_layers, cfg0.d_model, cfg0.n_heads, cfg0.n_kv_heads, cfg0.head_dim, cfg0.ffn, cfg0.vocab)
    assert (cfg.parallel, cfg.shared_ln, cfg.rotary_dim, cfg.hd_model, cfg.head_bias, cfg.act, cfg.gated) == \
           (cfg0.parallel, cfg0.shared_ln, cfg0.rotary_dim, cfg0.hd_model, cfg0.head_bias, cfg0.act, cfg0.gated)
    b, e = 1, cfg.n_layers
    got = loader.build_stage_weights_from_checkpoint(str(tmp_path), b, e, False, True, device="cpu")
    want = weights.build_stage_weights(cfg0, b, e, False, True, seed, "cpu")
    assert len(got["layers"]) == e - b
    for lg, lw in zip(got["layers"], want["layers"]):
        _same(lg, lw)
    for k in ("lnf_g", "lnf_b", "lm_head", "lm_head_b"):
        if k in want:
            assert torch.equal(got[k], want[k]), k
    first = loader.build_stage_weights_from_checkpoint(str(tmp_path), 0, 1, True, False, device="cpu")
    want0 = weights.build_stage_weights(cfg0, 0, 1, True, False, seed, "cpu")
    assert torch.equal(first["embed"], want0["embed"])
    if "pos_embed" in want0:
        assert torch.equal(first["pos_embed"], want0["pos_embed"])
    _same(first["layers"][0], want0["layers"][0])


def test_directory_without_weights_is_an_error(tmp_path):
    cfg = b200stage.get_config("tiny-llama-test")
    make_golden.hf_model(cfg, 1).config.save_pretrained(tmp_path)            # config.json only
    assert b200stage.get_config(str(tmp_path)).d_model == cfg.d_model
    with pytest.raises(FileNotFoundError):
        loader.build_stage_weights_from_checkpoint(str(tmp_path), 0, 1, True, False, device="cpu")
    with pytest.raises(FileNotFoundError):
        loader.calibrate(str(tmp_path), [np.arange(8)])


def _fp32_hidden(cfg, layers32, emb, ids):
    cos, sin = weights.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
    h = emb["embed"].to(torch.float16).to(torch.float32)[torch.from_numpy(ids)]
    for w in layers32:
        w = {k: (v.to(torch.float16).to(torch.float32) if v.dim() == 1 else v) for k, v in w.items()}
        h = loader._layer_fp32(cfg, w, h, cos, sin, {})
    return h.numpy()


def test_calibrated_smoothing_reduces_the_w8a8_error_on_an_outlier_channel_model(tmp_path):
    """A model with a few large-gain norm channels (what real LLMs show and SmoothQuant exists for): per-token int8
    activations lose the small channels next to the outliers.  Smoothing with CALIBRATED activation maxima
    (alpha = 0.5) must bring the W8A8 hidden state closer to the fp32 one than weight-only statistics do, and
    the statistics must survive the round trip through the file Stage.load picks up."""
    cfg = b200stage.get_config("tiny-gqa-test")
    seed = 11
    layers32 = [weights.add_outlier_gains(weights.init_layer_fp32(cfg, li, seed), cfg.d_model, li) for li in range(cfg.n_layers)]
    emb = weights.init_embed_fp32(cfg, seed)
    r = np.random.default_rng(0)
    calib = [r.integers(0, cfg.vocab, 48) for _ in range(4)]
    test_ids = r.integers(0, cfg.vocab, 64)
    stats = loader.calibrate(cfg, calib, seed, layer_fn=lambda li: layers32[li])
    assert len(stats) == cfg.n_layers and set(stats[0]) == {"qkv", "o", "gu", "d"}
    out_ch = weights.outlier_gain_channels(cfg.d_model, 0, 0)
    assert float(stats[0]["qkv"][out_ch].min()) > 4 * float(stats[0]["qkv"].median())      # the outliers were seen
    ref = _fp32_hidden(cfg, layers32, emb, test_ids)
    errs = {}
    for name, st in (("weight_only", None), ("calibrated", stats)):
        sw = dict(first=True, last=False, quant="w8a8", embed=emb["embed"].to(torch.float16),
                  layers=[weights.quantize_layer(cfg, w, None if st is None else st[li]) for li, w in enumerate(layers32)])
        cos, sin = O.rope_table(cfg.head_dim, cfg.max_pos, cfg.rope_theta, cfg.rope_scaling)
        cache = O.KVCacheOracle(cfg.n_layers, 1, cfg.n_kv_heads, 64, cfg.head_dim)
        got = O.stage_forward(test_ids, weights.to_numpy(sw), cfg.kernel_cfg(), cache, np.zeros(64, np.int64), np.arange(64), cos, sin)
        errs[name] = float(np.linalg.norm(got.astype(np.float64) - ref) / np.linalg.norm(ref))
    print(errs)
    assert errs["calibrated"] < 0.8 * errs["weight_only"], errs
    path = loader.save_stats(str(tmp_path), stats)
    back = loader.load_stats(str(tmp_path))
    assert os.path.basename(path) == loader.STATS_FILE and len(back) == len(stats)
    assert all(torch.equal(back[li][k], stats[li][k]) for li in range(len(stats)) for k in stats[li])


def test_fast_builder_calibration_is_the_same_on_every_rank():
    """bench.py calibrates while the device-side builder draws the layers.  A stage that starts at layer b re-draws and
    forwards layers 0..b-1 to get its input activations: its layers must come out identical to the same layers of a
    single-stage build (otherwise an N-stage pipeline would not decode the tokens of one GPU), and they must differ
    from the uncalibrated build (the statistics are used)."""
    cfg = b200stage.get_config("tiny-gqa-test")
    ids = np.random.default_rng(1).integers(0, cfg.vocab, 40)
    full = weights.build_stage_weights_fast(cfg, 0, cfg.n_layers, True, True, 5, "cpu", calib_ids=ids)
    part = weights.build_stage_weights_fast(cfg, 2, cfg.n_layers, False, True, 5, "cpu", calib_ids=ids)
    plain = weights.build_stage_weights_fast(cfg, 0, cfg.n_layers, True, True, 5, "cpu")
    assert len(part["layers"]) == cfg.n_layers - 2
    for a, b in zip(part["layers"], full["layers"][2:]):
        _same(a, b)
    assert set(full["layers"][0]) == set(plain["layers"][0])
    assert torch.equal(full["layers"][0]["ln1_g"], plain["layers"][0]["ln1_g"])
    assert not torch.equal(full["layers"][1]["sm_gu"], plain["layers"][1]["sm_gu"])
    assert torch.equal(full["lm_head"], plain["lm_head"])


@pytest.mark.parametrize("model", ["tiny-gqa-test", "tiny-neox-test"])
def test_head_rows_for_a_middle_stage(tmp_path, model):
    """A stage that is neither first nor last can hold a slice of the lm_head (the head shared by the stages of a
    deep pipeline): the rows it draws / reads are exactly the last stage's."""
    cfg = b200stage.get_config(model)
    seed, lo, hi = 3, 8, cfg.vocab - 16
    last = weights.build_stage_weights(cfg, cfg.n_layers - 1, cfg.n_layers, False, True, seed, "cpu")
    assert torch.equal(weights.head_rows_fp16(cfg, seed, lo, hi, "cpu", fast=False), last["lm_head"][lo:hi])
    make_golden.hf_for(cfg, seed).save_pretrained(tmp_path, safe_serialization=True)
    assert torch.equal(loader.load_head_rows(str(tmp_path), lo, hi, device="cpu"), last["lm_head"][lo:hi])


def test_fast_builder_head_rows_match_its_last_stage():
    cfg = b200stage.get_config("tiny-llama-test")
    last = weights.build_stage_weights_fast(cfg, cfg.n_layers - 1, cfg.n_layers, False, True, 5, "cpu")
    first = weights.build_stage_weights_fast(cfg, 0, 1, True, False, 5, "cpu")
    rows = weights.head_rows_fp16(cfg, 5, 40, 200, "cpu", fast=True)
    assert torch.equal(rows, last["lm_head"][40:200]) and torch.equal(rows, first["embed"][40:200])
