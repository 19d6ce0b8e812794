"""Checkpoint layout (SURVEY work item 10): {'net','acc','epoch'} both directions + wire codec."""
import base64
import io

import torch

from fedb200 import models
from fedb200.parallel.arena import Arena
from fedb200.utils import checkpoint as ckpt
from conftest import reference_src


def test_roundtrip_and_layout(tmp_path):
    m = models.MobileNet()
    a = Arena.adopt(m)
    p = tmp_path / "checkpoint" / "localhost:50051.pth"
    ckpt.save_checkpoint(str(p), a.state_dict(), acc=1, epoch=1)
    raw = torch.load(str(p), weights_only=True)           # what the reference does (torch.load)
    assert set(raw) == {"net", "acc", "epoch"} and raw["acc"] == 1 and raw["epoch"] == 1
    assert list(raw["net"].keys()) == list(models.MobileNet().state_dict().keys())
    assert raw["net"]["bn1.num_batches_tracked"].dtype == torch.int64
    fresh = models.MobileNet()
    fresh.load_state_dict(raw["net"])                      # strict load in a plain module


def test_wire_codec_is_reference_compatible():
    m = models.LeNet()
    raw = ckpt.dumps(m.state_dict())
    wire = ckpt.encode_wire(raw)
    assert wire[:4] == b"UEsD" and wire == base64.b64encode(raw)
    back = ckpt.load_checkpoint(ckpt.decode_wire(wire.decode("ascii")))
    assert all(torch.equal(v, back["net"][k]) for k, v in m.state_dict().items())


def test_reference_checkpoint_loads(ref_models):
    torch.manual_seed(3)
    ref = ref_models.MobileNet()
    buf = io.BytesIO()
    torch.save({"net": ref.state_dict(), "acc": 1, "epoch": 1}, buf)   # reference src/main.py:160-165
    state = ckpt.load_checkpoint(buf.getvalue())
    mine = models.MobileNet()
    a = Arena.adopt(mine)
    a.load_state_dict(state["net"])
    assert all(torch.equal(v, a.view(k)) for k, v in ref.state_dict().items())
    # and back: a file we write loads strictly into the reference module
    ref2 = ref_models.MobileNet()
    ref2.load_state_dict(ckpt.load_checkpoint(ckpt.dumps(a.state_dict()))["net"])


def test_fractional_counter_truncates_on_load():
    m = models.MobileNet()
    sd = m.state_dict()
    sd["bn1.num_batches_tracked"] = torch.tensor(195.5)   # what the reference coordinator saves
    a = Arena.adopt(models.MobileNet())
    a.load_state_dict(sd)
    assert int(a.view("bn1.num_batches_tracked")) == 195
