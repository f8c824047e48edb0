"""CPU: the oracle restatements against the golden vectors produced by the reference's own Python
(oracle/make_golden.py).  Bit-exact unless stated."""
import os

import numpy as np
import pytest
import torch

from helpers import golden, tiny_yolo_from_golden, vgg_block_from_golden
from oracle import binning as ob
from oracle import net_restated as onr


def _unpack(bits, shape):
    n = int(np.prod(shape))
    return np.unpackbits(bits)[:n].reshape(shape).astype(np.float32)


def test_binning_st_matches_reference_parse_data():
    g = golden("binning_st.npz")
    T, H, W = int(g["T"]), int(g["H"]), int(g["W"])
    rec = g["records"]
    # the reference loads events with t < load_until (psee_loader.load_delta_t) before binning
    rec = rec[rec[:, 0] < int(g["load_until_us"])]
    f = ob.bin_st(rec, int(g["first_event_time_us"]), int(g["step_us"]), T, H, W)
    ref = _unpack(g["features_bits"], (T, 2, H, W))
    assert int(ref.sum()) == int(g["n_ones"])
    assert np.array_equal(f, ref)
    # batched contract of the CUDA entry point (drops t >= t0 + T*step itself)
    fb = ob.bin_batch([g["records"]], [int(g["first_event_time_us"])], int(g["step_us"]), T, H, W, clip_x=True)
    assert np.array_equal(fb[:, 0], ref)


def test_binning_mt_matches_reference_parse_data():
    g = golden("binning_mt.npz")
    T, H, W, step = int(g["T"]), int(g["H"]), int(g["W"]), int(g["step_us"])
    ref = _unpack(g["features_bits"], (2, T, 2, H, W))
    rec = g["records"]
    for win in range(2):
        lo, hi = win * T * step, (win + 1) * T * step
        sl = rec[(rec[:, 0] >= lo) & (rec[:, 0] < hi)]
        assert np.array_equal(ob.bin_mt(sl, win * T, step, T, H, W), ref[win])
        fb = ob.bin_batch([rec], [lo], step, T, H, W, clip_x=False)
        assert np.array_equal(fb[:, 0], ref[win])


def test_dat_roundtrip_and_decode():
    rng = np.random.default_rng(5)
    rec = ob.make_records(rng, 500, 10, 5000, 240, 304)
    back, ev_type, ev_size, hw = ob.split_dat(ob.dat_bytes(rec, 240, 304))
    assert (ev_type, ev_size, hw) == (0, 8, (240, 304))
    assert np.array_equal(back, rec)
    t, x, y, p = ob.decode_events(rec)
    assert x.max() < 304 and y.max() < 240 and set(np.unique(p)) <= {0, 1}
    assert np.all(np.diff(t) >= 0)


def test_oracle_executor_matches_reference_blockgen_vgg():
    block, g = vgg_block_from_golden()
    X = torch.from_numpy(g["X"]).float()
    Y, state = onr.seq_block_forward(block, X)
    assert torch.equal(Y, torch.from_numpy(g["Y"]))
    last = [s for s in state[0] if s is not None][-1]
    assert torch.equal(last.v, torch.from_numpy(g["v_final"]))
    assert torch.equal(last.i, torch.from_numpy(g["i_final"]))
    assert 0.02 < float(Y.mean()) < 0.5  # not a dead net


def test_oracle_executor_matches_reference_tiny_yolo():
    model, X, g, digest_ok = tiny_yolo_from_golden()
    assert [k for k in model.state_dict().keys()] == [str(k) for k in g["keys"]], "state_dict keys differ from the reference's"
    assert digest_ok, "seeded weights differ from the reference's (torch RNG / construction order changed)"
    (anchors, cls, bbox), _ = onr.soda_forward(model, X)
    assert torch.equal(anchors, torch.from_numpy(g["anchors"]))
    assert torch.equal(cls, torch.from_numpy(g["cls"]))
    assert torch.equal(bbox, torch.from_numpy(g["bbox"]))
    assert float(np.min(g["rates"])) > 0.03


def test_box_utils_match_reference():
    from snn_for_object_detection_b200.utils import box
    from snn_for_object_detection_b200.utils.anchors import AnchorGenerator

    g = golden("box.npz")
    gen = AnchorGenerator(torch.from_numpy(g["sizes"]), torch.from_numpy(g["ratios"]))
    anchors = gen(torch.zeros(1, 4, int(g["fh"]), int(g["fw"])))
    assert torch.equal(anchors, torch.from_numpy(g["anchors"]))
    det = box.multibox_detection(torch.from_numpy(g["cls_probs"]), torch.from_numpy(g["offsets"]), anchors)
    assert torch.equal(det, torch.from_numpy(g["det"]))


def test_norse_restatement_known_answers():
    """PARITY UNPINNED (norse absent): hand-computed known answers of the published update rule."""
    from oracle import norse_restated as nr

    cell = nr.LIFCell()
    x = torch.tensor([6.0, 0.0])
    z, s = cell(x)  # i=6 -> v = 0.1*6 = 0.6 ; i' = 4.8
    assert torch.equal(z, torch.zeros(2))
    assert torch.allclose(s.v, torch.tensor([0.6, 0.0])) and torch.allclose(s.i, torch.tensor([4.8, 0.0]))
    z, s = cell(torch.zeros(2), s)  # v = 0.6 + 0.1*(4.8-0.6) = 1.02 > 1 -> spike, reset
    assert torch.equal(z, torch.tensor([1.0, 0.0]))
    assert float(s.v[0]) == 0.0 and abs(float(s.i[0]) - 3.84) < 1e-6
    li = nr.LICell()
    v, s = li(torch.tensor([1.0]))
    assert abs(float(v) - 0.1) < 1e-7 and abs(float(s.i) - 0.8) < 1e-7
    # surrogate gradient: 1 / (100*|v-1| + 1)^2
    xg = torch.tensor([0.5], requires_grad=True)
    nr.threshold(xg, "super", 100.0).backward()
    assert abs(float(xg.grad) - 1.0 / (100 * 0.5 + 1) ** 2) < 1e-9


def test_roi_restatement_matches_reference_golden():
    """oracle/roi_restated.py against the output of the reference's OWN utils/roi.py (tests/golden/roi.npz):
    anchor -> box map (mask, classes) identical, incl. the padded -1 rows that claim anchors; offsets to 1 ulp of
    np.log vs torch.log."""
    from oracle import roi_restated as rr

    g = golden("roi.npz")
    off, msk, cls = rr.roi(g["anchors"], g["labels"], float(g["iou_threshold"]))
    assert np.array_equal(cls, g["class_labels"])
    assert np.array_equal(msk, g["bbox_mask"])
    assert int((g["class_labels"] > 0).sum()) > 100 and int(g["bbox_mask"][2, :, 0].sum()) == 6
    np.testing.assert_allclose(off, g["bbox_offset"], rtol=1e-6, atol=1e-6)


def test_roi_restatement_live_reference():
    """Same check against the live reference on fresh random boxes, when /root/reference is present (build container)."""
    from oracle import ref_shim
    from oracle import roi_restated as rr

    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    ref = ref_shim.load_reference()
    rng = np.random.default_rng(5)
    ctr = rng.random((700, 2), dtype=np.float32)
    wh = rng.random((700, 2), dtype=np.float32) * 0.3 + 0.05
    anchors = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=1)
    labels = np.full((4, 9, 5), -1.0, np.float32)
    for b in range(4):
        for k in range(2 * b + 1):
            labels[b, k, 0] = k % 3
            labels[b, k, 1:] = anchors[rng.integers(700)] + (rng.random(4, dtype=np.float32) - 0.5) * 0.1
    want = ref.roi.RoI(0.4)(torch.from_numpy(anchors), torch.from_numpy(labels.copy()))
    off, msk, cls = rr.roi(anchors, labels, 0.4)
    assert np.array_equal(cls, want[2].numpy()) and np.array_equal(msk, want[1].numpy())
    np.testing.assert_allclose(off, want[0].numpy(), rtol=1e-6, atol=1e-6)


def test_sli_restatement_matches_reference_golden():
    """oracle/sli_restated.py against outputs of the reference's OWN sli_feed_forward_step (models/modules/sli.py),
    bit for bit - this neuron's arithmetic is in the reference tree, so unlike LIF / LI it is pinned."""
    from oracle import sli_restated as sr

    g = golden("sli.npz")
    out, v, i = sr.sli_scan(torch.from_numpy(g["x"]))
    assert torch.equal(out, torch.from_numpy(g["out"]))
    assert torch.equal(v, torch.from_numpy(g["v"])) and torch.equal(i, torch.from_numpy(g["i"]))
    assert float((out.abs() > 1).float().mean()) > 0.2   # the saturation branch is exercised


def test_synapse_restatement_matches_reference_golden():
    """oracle/synapse_restated.py against outputs of the reference's OWN synapse_feed_forward_step
    (models/modules/synapse.py), bit for bit, with and without synaptic inhibition."""
    from oracle import synapse_restated as sr

    g = golden("synapse.npz")
    for tag, sigma in (("s0", 0.0), ("s07", 0.7)):
        out, p = sr.synapse_scan(torch.from_numpy(g["x"]), sigma_inhibition=sigma)
        assert torch.equal(out, torch.from_numpy(g["out_" + tag])) and torch.equal(p, torch.from_numpy(g["p_" + tag]))
        assert float(out.mean()) > 0.01


def test_st_sample_stream_window_selection_matches_reference_golden():
    """events.STSampleStream.next_window against the reference's own STPropheseeDataset.parse_data golden
    (binning_st.npz): same window start, same records, oracle binning of that window == the reference's features."""
    from snn_for_object_detection_b200.events import STSampleStream

    g = golden("binning_st.npz")
    T, H, W, step_us = int(g["T"]), int(g["H"]), int(g["W"]), int(g["step_us"])
    dat = ob.dat_bytes(g["records"], H, W)
    st = STSampleStream([dat], [[6]], num_steps=T, time_shift=1, time_step_ms=step_us // 1000, height=H, width=W,
                        device="cpu")
    win, alive = st.next_window(0)
    assert alive and win is not None
    rec, t0, label_step = win
    assert t0 == int(g["first_event_time_us"]) and label_step == 6
    assert st.readers[0].current_time == int(g["load_until_us"])
    want = np.unpackbits(g["features_bits"])[: T * 2 * H * W].reshape(T, 2, H, W).astype(np.float32)
    assert np.array_equal(ob.bin_st(np.ascontiguousarray(rec), t0, step_us, T, H, W), want)
    assert int(rec[0, 0]) >= t0 and int(rec[-1, 0]) < int(g["load_until_us"])
    # no label far enough ahead any more -> recording dropped; a sparse window -> skipped but alive
    assert st.next_window(0) == (None, False)
    st2 = STSampleStream([dat], [[6]], num_steps=T, time_shift=1, time_step_ms=step_us // 1000, height=H, width=W,
                         events_threshold=10 ** 6, device="cpu")
    assert st2.next_window(0) == (None, True)


def test_st_sample_stream_matches_live_reference_parse_data():
    """Several recordings, several labels each: the sequence of (accepted / skipped / dropped) decisions and every
    accepted window's features equal the live reference's parse_data (build container only)."""
    import tempfile

    from oracle import ref_shim
    from snn_for_object_detection_b200.events import STSampleStream

    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    ref = ref_shim.load_reference()
    H, W, T, step_ms, shift = 60, 80, 4, 16, 1
    step_us = step_ms * 1000
    rng = np.random.default_rng(77)
    recs = [ob.make_records(rng, n, 0, 40 * step_us, H, W) for n in (90000, 30000)]
    labels = [[6, 7, 15, 16, 30], [5, 22]]
    with tempfile.TemporaryDirectory() as d:
        paths = []
        for k, r in enumerate(recs):
            paths.append(os.path.join(d, f"r{k}_td.dat"))
            with open(paths[-1], "wb") as f:
                f.write(ob.dat_bytes(r, H, W))
        ds = ref.datasets.STPropheseeDataset(num_steps=T, time_shift=shift, gt_files=[], data_files=[], time_step=step_ms,
                                             num_load_file=1, name="gen1")
        ds._height, ds._width = H, W
        ds.events_threshold = 1500
        st = STSampleStream(paths, labels, T, shift, step_ms, H, W, events_threshold=1500, device="cpu")
        for k in range(2):
            gt = torch.tensor([[float(s), 0.0, 0.1, 0.1, 0.6, 0.7] for s in labels[k]])
            loader = ref.psee.PSEELoader(paths[k])
            for _ in range(8):
                data, ok = ds.parse_data(gt, loader)
                win, alive = st.next_window(k)
                assert alive == ok and (data is None) == (win is None)
                if win is not None:
                    rec, t0, _ = win
                    assert np.array_equal(ob.bin_st(np.ascontiguousarray(rec), t0, step_us, T, H, W), data[0].numpy())
                if not ok:
                    break


def _scan_cell(step, cell_state, x):
    vs, cs, state = [], [], cell_state
    for t in range(x.shape[0]):
        out, state = step(x[t], state)
        vs.append(state.v)
        cs.append(state.i)
    return torch.stack(vs), torch.stack(cs), out


def test_li_and_subthreshold_lif_pinned_to_the_reference_sli_cell():
    """Rows a8/a9: ``tests/golden/li_from_sli.npz`` holds T = 16 steps of the reference's OWN
    ``sli_feed_forward_step`` (models/modules/sli.py:110-126) with ``v_st = 1e4``, where its saturation factor is
    exactly 1.0f and the function is the LI update.  The norse restatement must reproduce it bit for bit: LI on
    large inputs, and LIF on inputs that never reach the threshold (no spike, no reset => same (v, i))."""
    from oracle import norse_restated as nr

    g = golden("li_from_sli.npz")
    with torch.no_grad():
        x = torch.from_numpy(g["x_big"])
        s0 = nr.LIState(v=torch.zeros_like(x[0]), i=torch.zeros_like(x[0]))
        v, i, _ = _scan_cell(lambda a, s: nr.li_feed_forward_step(a, s), s0, x)
        assert torch.equal(v, torch.from_numpy(g["v_big"])) and torch.equal(i, torch.from_numpy(g["i_big"]))
        x = torch.from_numpy(g["x_sub"])
        s0 = nr.LIFFeedForwardState(v=torch.zeros_like(x[0]), i=torch.zeros_like(x[0]))
        zs = []

        def lif(a, s):
            z, s = nr.lif_feed_forward_step(a, s)
            zs.append(z)
            return z, s

        v, i, _ = _scan_cell(lif, s0, x)
        assert float(torch.stack(zs).abs().max()) == 0.0
        assert torch.equal(v, torch.from_numpy(g["v_sub"])) and torch.equal(i, torch.from_numpy(g["i_sub"]))
        # the cells' own initial state (v = v_leak = 0, i = 0) gives the same sequence
        out, st = nr.LICell()(torch.from_numpy(g["x_big"])[0])
        assert torch.equal(out, torch.from_numpy(g["v_big"])[0])


def test_li_pin_live_reference():
    """Same check against the live reference function on fresh inputs (build container only)."""
    import sys

    from oracle import norse_restated as nr
    from oracle import ref_shim

    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    ref_shim.load_reference()
    sli = sys.modules["models.modules.sli"]
    prm = sli.SLIParameters(v_st=torch.as_tensor(1.0e4))
    torch.manual_seed(99)
    x = torch.randn(20, 7, 33) * 6.0
    with torch.no_grad():
        sr = sli.SLIState(v=torch.zeros_like(x[0]), i=torch.zeros_like(x[0]))
        so = nr.LIState(v=torch.zeros_like(x[0]), i=torch.zeros_like(x[0]))
        for t in range(x.shape[0]):
            vr, sr = sli.sli_feed_forward_step(x[t], sr, prm, 0.001)
            vo, so = nr.li_feed_forward_step(x[t], so)
            assert torch.equal(vr, vo) and torch.equal(sr.i, so.i)


def test_pure_torch_containers_mirror_the_product_module_tree():
    """oracle/containers.py (used by bench.py's CPU arm so that it never imports the product) builds the same
    module tree: identical state_dict keys and shapes, and the oracle executor gives identical outputs on either
    container with the same weights."""
    import snn_for_object_detection_b200 as pkg
    from oracle import containers as oc
    from oracle import net_restated as onr

    torch.manual_seed(3)
    prod = pkg.TinyYolo(num_classes=2).eval()
    cont = oc.TinyYolo(num_classes=2).eval()
    sp, sc = prod.state_dict(), cont.state_dict()
    assert list(sp.keys()) == list(sc.keys())
    assert all(sp[k].shape == sc[k].shape for k in sp)
    cont.load_state_dict(sp, strict=True)
    X = (torch.rand(2, 1, 2, 64, 64) < 0.2).float()
    for m in list(prod.modules()) + list(cont.modules()):
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.fill_(5.0)
    (a0, c0, b0), _ = onr.soda_forward(prod, X)
    (a1, c1, b1), _ = onr.soda_forward(cont, X)
    assert torch.equal(a0, a1) and torch.equal(c0, c1) and torch.equal(b0, b1)
    from helpers import vgg_cfg

    vp, vc = pkg.BlockGen(2, vgg_cfg()), oc.vgg_net()
    assert list(vp.state_dict().keys()) == list(vc.state_dict().keys())
    vc.load_state_dict(vp.state_dict(), strict=True)
    Xv = (torch.rand(3, 1, 2, 24, 32) < 0.3).float()
    y0, _ = onr.seq_block_forward(vp.eval(), Xv)
    y1, _ = onr.seq_block_forward(vc.eval(), Xv)
    assert torch.equal(y0, y1)
