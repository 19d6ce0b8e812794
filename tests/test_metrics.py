"""Observability utilities: the reference's progress bar and time formatter (src/utils.py:51-124) without
its import-time `stty size` dependency, plus the JSONL sink that replaces its prints."""
import io
import json

from fedb200.utils.metrics import JsonlSink, ProgressBar, format_time, nvtx_range


def test_format_time_two_most_significant_units():
    assert format_time(0) == "0ms" and format_time(0.012) == "12ms" and format_time(3.25) == "3s250ms"
    assert format_time(72) == "1m12s" and format_time(3600 * 5 + 61) == "5h1m" and format_time(86400 * 2 + 3600) == "2D1h"


def test_progress_bar_is_silent_without_a_tty_and_draws_when_forced():
    quiet = io.StringIO()
    bar = ProgressBar(4, stream=quiet)                   # StringIO is not a TTY: nothing is written (headless boxes, logs)
    line = bar.update(0, "Loss: 2.303 | Acc: 10.000% (12/128)")
    assert quiet.getvalue() == "" and "1/4" in line and "Loss: 2.303" in line and "Step:" in line and "Tot:" in line
    loud = io.StringIO()
    bar = ProgressBar(2, width=10, stream=loud, force=True)
    bar.update(0)
    bar.update(1, "done")
    text = loud.getvalue()
    assert text.count("\r") == 1 and text.endswith("\n") and "[=========>]" in text and "2/2" in text


def test_jsonl_sink_keeps_and_writes_records(tmp_path):
    path = tmp_path / "m.jsonl"
    sink = JsonlSink(str(path))
    rec = sink.write("round", round=3, live=2, total_s=0.25)
    sink.write("eval", acc=0.5)
    lines = [json.loads(x) for x in path.read_text().splitlines()]
    assert [x["event"] for x in lines] == ["round", "eval"] and lines[0]["round"] == 3 and rec["live"] == 2
    assert len(sink.records) == 2 and "ts" in lines[1]
    assert JsonlSink(None).write("x", a=1)["a"] == 1     # no file: records only


def test_nvtx_range_is_a_no_op_without_cuda():
    with nvtx_range("phase"):
        pass


def test_device_phase_is_a_no_op_without_a_timer():
    from fedb200.utils.metrics import device_phase
    ran = []
    with device_phase(None, "agg"):
        ran.append(1)
    assert ran == [1]
