"""obst/run/summary.py: the reference's scalar names (src/run/run.py:123-132) and a TensorBoard event file written by
hand -- read back by TensorBoard's own loader / accumulator and compared, field by field, with what TensorBoard's own
TF2-style writer (`tensorboard.summary.Writer.add_scalar`) puts into an event for the same scalar.  CPU only."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))

event_file_loader = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
event_accumulator = pytest.importorskip("tensorboard.backend.event_processing.event_accumulator")
tensor_util = pytest.importorskip("tensorboard.util.tensor_util")
tb_summary = pytest.importorskip("tensorboard.summary")


def _loss(nll_sum, z_sum, hits, n, z):
    from obst.model import Loss
    return Loss(torch.tensor([nll_sum, z_sum, hits], dtype=torch.float64), n, z)


def test_log_dict_carries_the_reference_names():
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.run.summary import log_dict
    from oracle import obst_oracle as orc
    params = ModelParameter(gpt_mode(dict(orc.TINY_CONFIG)))
    total, first, last = _loss(500.0, 3000.0, 10.0, 100, 1e-4), _loss(260.0, 1500.0, 4.0, 50, 1e-4), _loss(
        240.0, 1500.0, 6.0, 50, 1e-4)
    out = (None, None, torch.tensor(0.005), total, first, last, None, last, last, [], {})
    d = log_dict(params, out)
    assert list(d) == ["learning_rate", "first_loss", "last_loss", "mean_loss", "token_loss", "accuracy"]
    assert abs(d["learning_rate"] - 0.005) < 1e-9
    assert abs(d["mean_loss"] - (5.0 + 1e-4 * 30.0)) < 1e-12
    assert abs(d["first_loss"] - (5.2 + 1e-4 * 30.0)) < 1e-12
    assert abs(d["last_loss"] - (4.8 + 1e-4 * 30.0)) < 1e-12 and d["token_loss"] == d["last_loss"]
    assert abs(d["accuracy"] - 0.12) < 1e-12
    assert list(log_dict(params, out[:8] + (None,) + out[9:])) == ["learning_rate", "first_loss", "last_loss",
                                                                   "mean_loss", "token_loss"]


def test_event_file_is_read_by_tensorboard(tmp_path):
    from obst.run.summary import SummaryWriter
    steps = {0: {"learning_rate": 0.0, "mean_loss": 5.55, "accuracy": 0.004},
             4096: {"learning_rate": 0.01, "mean_loss": 3.25, "accuracy": 0.31},
             1 << 40: {"learning_rate": 1e-9, "mean_loss": -2.5, "accuracy": 1.0}}
    with SummaryWriter(str(tmp_path / "mine")) as w:
        for step, values in steps.items():
            w.scalars(values, step)
        w.flush()
        path = w.path
    events = list(event_file_loader.EventFileLoader(path).Load())
    assert events[0].file_version == "brain.Event:2"
    got = {}
    for e in events[1:]:
        assert len(e.summary.value) == 1
        v = e.summary.value[0]
        got.setdefault(e.step, {})[v.tag] = float(tensor_util.make_ndarray(v.tensor))
        assert v.tensor.dtype == 1 and len(v.tensor.tensor_shape.dim) == 0
    assert {s: {k: pytest.approx(x, rel=1e-6) for k, x in d.items()} for s, d in steps.items()} == got
    # the accumulator (what the TensorBoard UI is fed from) sees three tags with three points each
    acc = event_accumulator.EventAccumulator(str(tmp_path / "mine"), size_guidance={event_accumulator.TENSORS: 0})
    acc.Reload()
    tags = acc.Tags()
    assert sorted(tags["tensors"]) == ["accuracy", "learning_rate", "mean_loss"] or \
        sorted(tags["scalars"]) == ["accuracy", "learning_rate", "mean_loss"], tags
    kind = "tensors" if tags["tensors"] else "scalars"
    points = acc.Tensors("mean_loss") if kind == "tensors" else acc.Scalars("mean_loss")
    assert [p.step for p in points] == list(steps)


def test_events_match_tensorboards_own_tf2_style_writer(tmp_path):
    """Same scalar through `tensorboard.summary.Writer.add_scalar` (which writes what tf.summary.scalar writes): the
    value protos agree in tag, tensor dtype / shape / value and plugin metadata."""
    from obst.run.summary import SummaryWriter
    theirs_dir = str(tmp_path / "theirs")
    w = tb_summary.Writer(theirs_dir)
    w.add_scalar("mean_loss", 3.25, step=7)
    w.close()
    theirs_file = [f for f in os.listdir(theirs_dir) if "tfevents" in f]
    assert len(theirs_file) == 1
    theirs = [e for e in event_file_loader.EventFileLoader(os.path.join(theirs_dir, theirs_file[0])).Load()
              if e.summary.value]
    with SummaryWriter(str(tmp_path / "mine")) as mine_writer:
        mine_writer.scalar("mean_loss", 3.25, 7)
        mine_writer.scalar("mean_loss", 3.5, 8)
        path = mine_writer.path
    mine = [e for e in event_file_loader.EventFileLoader(path).Load() if e.summary.value]
    a, b = theirs[0], mine[0]
    assert a.step == b.step == 7
    va, vb = a.summary.value[0], b.summary.value[0]
    assert va.tag == vb.tag
    assert va.metadata.plugin_data.plugin_name == vb.metadata.plugin_data.plugin_name == "scalars"
    assert va.metadata.data_class == vb.metadata.data_class
    assert va.tensor.dtype == vb.tensor.dtype
    assert list(va.tensor.tensor_shape.dim) == list(vb.tensor.tensor_shape.dim) == []
    assert float(tensor_util.make_ndarray(va.tensor)) == float(tensor_util.make_ndarray(vb.tensor)) == 3.25
    # like TensorFlow's writer, later events of a tag carry no metadata
    assert not mine[1].summary.value[0].HasField("metadata")
