"""The oracle (oracle/ddp_oracle.py) against the golden vectors produced by the real upstream routines
(oracle/make_golden.py). CPU only. An oracle that fails here cannot be trusted as a checker."""
import json
import os

import numpy as np
import torch

from oracle import ddp_oracle as O


def test_philox_known_answers(golden_dir):
    kat = json.load(open(os.path.join(golden_dir, "philox.json")))["kat"]
    for k in kat:
        got = O.philox4x32_10_raw(*[np.array([c]) for c in k["ctr"]], *k["key"])[0]
        assert [int(v) for v in got] == k["out"]


def test_dropout_mask_properties():
    keep = O.dropout_keep_mask(1 << 16, 0.5, seed=1234)
    assert abs(keep.mean() - 0.5) < 0.01
    assert (O.dropout_keep_mask(1 << 16, 0.5, seed=1234) == keep).all()
    assert (O.dropout_keep_mask(1 << 16, 0.5, seed=1235) != keep).any()
    assert O.dropout_keep_mask(100, 0.0, seed=1).all()


def test_sampler_matches_upstream(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "sampler.json")))
    import hashlib
    for c in g["cases"]:
        if c.get("shuffle", True):
            gen = torch.Generator().manual_seed(c["seed"] + c["epoch"])
            perm = torch.randperm(c["n"], generator=gen).tolist()
        else:
            perm = list(range(c["n"]))
        idx = O.sampler_indices(perm, c["world"], c["rank"], c["drop_last"])
        assert len(idx) == c["len"], c
        assert idx[:6].tolist() == c["first"], c
        assert hashlib.sha256(idx.tobytes()).hexdigest()[:16] == c["sha16"], c


def test_survey_sampler_hashes(golden_dir):
    """The hashes SURVEY.md 8(c) lists were produced independently of make_golden.py."""
    g = json.load(open(os.path.join(golden_dir, "sampler.json")))
    by = {(c["n"], c["world"], c["epoch"], c["rank"]): c["sha16"] for c in g["cases"] if c.get("shuffle", True)
          and not c["drop_last"]}
    assert by[(50000, 1, 0, 0)] == "62d58f141ce65ae7"
    assert by[(50000, 2, 0, 0)] == "4ebfccbb258c5eb7" and by[(50000, 2, 0, 1)] == "54083bebbcb31fa8"
    assert by[(50000, 8, 0, 0)] == "df736513afad78cf" and by[(50000, 8, 0, 7)] == "e7f548809fbcca57"
    assert by[(50000, 8, 1, 0)] == "0671688e082c7d55" and by[(50000, 8, 1, 7)] == "370c0016818732f1"
    assert by[(10000, 8, 0, 0)] == "7b7e639a3bf3b81e" and by[(10000, 2, 0, 1)] == "d227c99b4b779987"
    assert g["batch_counts_bs128"] == {"1": {"batches": 391, "last": 80}, "2": {"batches": 196, "last": 40},
                                       "4": {"batches": 98, "last": 84}, "8": {"batches": 49, "last": 106}}


def test_bucket_assignment_matches_upstream(golden_dir):
    for case in json.load(open(os.path.join(golden_dir, "buckets.json"))):
        assert O.bucket_assignment(case["nbytes"], case["caps"]) == case["buckets"]
    first = json.load(open(os.path.join(golden_dir, "buckets.json")))[0]
    sizes = [sum(first["nbytes"][i] for i in b) for b in first["buckets"]]
    assert sizes == [67289128, 151011328, 9878784]


def test_adam_matches_upstream(golden_dir):
    g = np.load(os.path.join(golden_dir, "adam.npz"))
    for wd in (0.0, 0.01):
        p, m, v = g["p0"], np.zeros_like(g["p0"]), np.zeros_like(g["p0"])
        for t in (1, 2, 3):
            p, m, v = O.adam_step(p, g[f"g{t}"], m, v, t, lr=1e-3, weight_decay=wd)
            tag = f"wd{wd}_t{t}"
            np.testing.assert_allclose(p, g[f"p_{tag}"], rtol=2e-6, atol=2e-7)
            np.testing.assert_allclose(m, g[f"m_{tag}"], rtol=2e-6, atol=1e-7)
            np.testing.assert_allclose(v, g[f"v_{tag}"], rtol=2e-6, atol=1e-9)


def test_clip_matches_upstream(golden_dir):
    g = np.load(os.path.join(golden_dir, "clip.npz"))
    grads = [g[f"g{i}"] for i in range(4)]
    for mn in (1.0, 100.0):
        out, total = O.clip_grad_norm(grads, mn)
        assert abs(total - float(g[f"norm_total_{mn}"])) < 1e-5 * total
        for i, o in enumerate(out):
            np.testing.assert_allclose(o, g[f"norm_{mn}_g{i}"], rtol=2e-6, atol=1e-7)
    for i, gr in enumerate(grads):
        np.testing.assert_array_equal(O.clip_grad_value(gr, 0.5), g[f"value_0.5_g{i}"])


def test_layers_match_upstream(golden_dir):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    y, gate = O.linear_fwd(g["lin_x"], g["lin_w"], g["lin_b"], relu=True)
    np.testing.assert_allclose(y, g["lin_y"], rtol=1e-5, atol=1e-5)
    dx, dw, db = O.linear_bwd(g["lin_dy"], gate, g["lin_x"], g["lin_w"])
    np.testing.assert_allclose(dx, g["lin_dx"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(dw, g["lin_dw"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(db, g["lin_db"], rtol=1e-5, atol=1e-5)
    loss, dz = O.softmax_ce(g["ce_z"], g["ce_labels"])
    assert abs(float(loss) - float(g["ce_loss"])) < 1e-6
    np.testing.assert_allclose(dz, g["ce_dz"], rtol=1e-5, atol=1e-7)
    yp, idx = O.relu_maxpool_fwd(g["pool_x"])
    np.testing.assert_array_equal(yp, g["pool_y"])
    np.testing.assert_allclose(O.relu_maxpool_bwd(g["pool_dy"], idx, g["pool_x"].shape), g["pool_dx"], rtol=0,
                               atol=1e-6)


def test_syncbn_matches_global_batch_bn(golden_dir):
    g = np.load(os.path.join(golden_dir, "layers.npz"))
    xs = [g["bn_x0"], g["bn_x1"]]
    ys, rm, rv, mean, invstd = O.syncbn_forward(xs, g["bn_w"], g["bn_b"], np.zeros(5), np.ones(5), relu=True)
    np.testing.assert_allclose(np.concatenate(ys), g["bn_y"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(rm, g["bn_rm"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(rv, g["bn_rv"], rtol=1e-5, atol=1e-6)
    # the per-rank stats + merge path must agree with the global statistics
    st = [O.bn_local_stats(x) for x in xs]
    m2, i2, _ = O.bn_merge_stats([s[0] for s in st], [s[1] for s in st], [x.shape[0] * 24 for x in xs])
    np.testing.assert_allclose(m2, mean, rtol=1e-10)
    np.testing.assert_allclose(i2, invstd, rtol=1e-9)
    dys = [g["bn_dy"][:3], g["bn_dy"][3:]]
    dxs, gws, gbs = O.syncbn_backward(dys, xs, g["bn_w"], mean, invstd, ys_relu=ys)
    np.testing.assert_allclose(np.concatenate(dxs), g["bn_dx"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(sum(gws), g["bn_gw"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(sum(gbs), g["bn_gb"], rtol=1e-4, atol=1e-5)


def test_evaluate_restatement_matches_reference_evaluate(golden_dir):
    """oracle.evaluate_batches against the totals the reference's own evaluate() returned (make_golden.gen_eval)."""
    g = np.load(os.path.join(golden_dir, "eval.npz"))
    bs = int(g["batch"])
    z, y = g["logits"], g["labels"]
    zs = [z[i:i + bs] for i in range(0, len(z), bs)]
    ys = [y[i:i + bs] for i in range(0, len(y), bs)]
    loss, correct, total = O.evaluate_batches(zs, ys)
    assert abs(loss - float(g["total_test_loss"])) < 1e-5 * abs(float(g["total_test_loss"]))
    assert correct == float(g["correct"]) == 10.0 and total == float(g["total"]) == 20.0
    # the head restatement reproduces the recorded logits from the recorded hidden activations
    z2, _ = O.linear_fwd(g["hidden"], g["w"], g["b"])
    np.testing.assert_allclose(z2, z, rtol=1e-4, atol=1e-5)


def test_input_pipeline_oracle_matches_torchvision(golden_dir):
    """oracle.input_pipeline (restatement of Pillow's 8-bit bilinear resample + ToTensor + Normalize) against the real
    torchvision transform's outputs: uint8 stage and float stage bit for bit."""
    import hashlib
    from oracle import input_pipeline as ip
    g = np.load(os.path.join(golden_dir, "input_pipeline.npz"))
    for i in range(g["images"].shape[0]):
        r = ip.resize_bilinear_u8(g["images"][i])
        assert np.array_equal(r, g["resized_u8"][i])
        out = ip.transform(g["images"][i], bool(g["flips"][i]))
        assert np.array_equal(out[:, :16, :], g["first_rows"][i])
        assert hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest() == str(g["sha256"][i])


def test_resize_coeffs_entry_point_matches_oracle():
    """b200ddp_resize_coeffs is host-only: the C restatement of Resample.c's coefficient set-up equals the oracle's."""
    import ctypes
    from b200ddp import _lib
    from oracle import input_pipeline as ip
    L = _lib.load()
    for n_in, n_out in [(32, 224), (32, 32), (17, 224), (48, 96), (1, 8)]:
        tab = (ctypes.c_int * (5 * n_out))()
        assert L.b200ddp_resize_coeffs(n_in, n_out, tab) == 0
        t = np.array(list(tab), np.int32)
        xm, cnt, co = ip.resample_coeffs(n_in, n_out)
        assert np.array_equal(t[:n_out], xm) and np.array_equal(t[n_out:2 * n_out], cnt)
        assert np.array_equal(t[2 * n_out:].reshape(n_out, 3), co[:, :3])
    tab = (ctypes.c_int * 40)()
    assert L.b200ddp_resize_coeffs(64, 8, tab) != 0          # reductions are refused, with a message
    assert b"enlargements" in L.b200ddp_last_error()
