"""Logger formats, event files (validated with the installed tensorboard) and crc32c
(SURVEY.md §4 item 1)."""
import io
import json
import os
import struct

import numpy as np
import pytest

from distributedtensorflow_b200 import logger
from distributedtensorflow_b200.utils import events, hostlib


def test_crc32c_known_vector_and_mask():
    assert events.crc32c_py(b"123456789") == 0xE3069283
    assert events.crc32c(b"123456789") == 0xE3069283
    lib = hostlib.lib()
    assert lib.dtf_crc32c(b"123456789", 9) == 0xE3069283
    assert lib.dtf_crc32c(b"", 0) == 0
    data = bytes(range(256)) * 37 + b"tail"
    assert lib.dtf_crc32c(data, len(data)) == events.crc32c_py(data)
    part = lib.dtf_crc32c(data[:100], 100)
    assert lib.dtf_crc32c_extend(part, data[100:], len(data) - 100) == events.crc32c_py(data)
    for c in (0, 1, 0xE3069283, 0xFFFFFFFF):
        m = events.mask_crc(c)
        assert m == lib.dtf_crc32c_mask(c) == ((((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF)
        assert events.unmask_crc(m) == c == lib.dtf_crc32c_unmask(m)


def test_masked_crc_matches_tensorboard():
    tb = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    for payload in (b"", b"a", b"hello world" * 100):
        assert events.masked_crc32c(payload) == tb.masked_crc32c(payload)


@pytest.mark.parametrize("native", [True, False])
def test_event_file_round_trip_through_tensorboard(tmp_path, native):
    w = events.EventsWriter(str(tmp_path / "events"), use_native=native)
    assert w.is_native == native
    assert os.path.basename(w.filename).startswith("events.out.tfevents.")
    w.write_scalars({"Global Step": 7, "Loss": 0.25}, 7, wall_time=1234.5)
    w.write_scalars({"Loss": 0.125}, 8)
    w.Flush()
    w.Close()
    evs = list(events.summary_iterator(w.filename))
    assert evs[0].file_version == "brain.Event:2"
    assert evs[1].step == 7 and evs[1].wall_time == 1234.5
    assert {v.tag: v.simple_value for v in evs[1].summary.value} == {"Global Step": 7.0, "Loss": 0.25}
    assert evs[2].step == 8 and evs[2].summary.value[0].simple_value == 0.125
    # the installed tensorboard is the independent oracle for the on-disk format
    loader_mod = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    tb_events = list(loader_mod.LegacyEventFileLoader(w.filename).Load())
    assert tb_events[0].file_version == "brain.Event:2"
    assert tb_events[1].step == 7 and abs(tb_events[1].wall_time - 1234.5) < 1e-9
    assert [(v.tag, v.simple_value) for v in tb_events[1].summary.value] == [("Global Step", 7.0), ("Loss", 0.25)]


def test_native_and_python_writers_are_byte_identical(tmp_path):
    a = events.EventsWriter(str(tmp_path / "a"), use_native=True)
    b = events.EventsWriter(str(tmp_path / "b"), use_native=False)
    for w in (a, b):
        w.write_scalars({"x": 1.5, "y": -2.0}, 3, wall_time=99.0)
        w.Close()
    ra = list(events.tf_record_iterator(a.filename))[1:]
    rb = list(events.tf_record_iterator(b.filename))[1:]
    assert ra == rb and len(ra) == 1


def test_tfrecord_detects_corruption(tmp_path):
    w = events.EventsWriter(str(tmp_path / "c"), use_native=False)
    w.write_scalars({"x": 1.0}, 1)
    w.Close()
    raw = bytearray(open(w.filename, "rb").read())
    raw[-6] ^= 0xFF
    open(w.filename, "wb").write(raw)
    with pytest.raises(IOError):
        list(events.tf_record_iterator(w.filename))
    lib = hostlib.lib()
    import ctypes as C
    r = lib.dtf_tfrecord_open(os.fsencode(w.filename))
    data, n = C.c_void_p(), C.c_size_t()
    assert lib.dtf_tfrecord_next(r, C.byref(data), C.byref(n)) == 1   # header record is intact
    assert lib.dtf_tfrecord_next(r, C.byref(data), C.byref(n)) < 0    # corrupted one is rejected
    lib.dtf_tfrecord_close(r)


def test_human_output_format_table_bytes():
    buf = io.StringIO()
    fmt = logger.HumanOutputFormat(buf)
    fmt.writekvs({"b": 2.5, "a": 3, "Zeta": "x" * 30})
    expected = ("----------------------------------\n"
                "| a    | 3                       |\n"
                "| b    | 2.5                     |\n"
                "| Zeta | xxxxxxxxxxxxxxxxxxxx... |\n"
                "----------------------------------\n")
    assert buf.getvalue() == expected
    buf2 = io.StringIO()
    logger.HumanOutputFormat(buf2).writeseq(["hello", "world"])
    assert buf2.getvalue() == "hello world\n"


def test_csv_header_rewrite_and_json_lines(tmp_path):
    c = logger.CSVOutputFormat(str(tmp_path / "p.csv"))
    c.writekvs({"a": 1, "b": 2})
    c.writekvs({"a": 3, "b": 4, "c": 5})
    c.writekvs({"c": 6})
    c.close()
    assert open(tmp_path / "p.csv").read() == "a,b,c\n1,2,\n3,4,5\n,,6\n"
    j = logger.JSONOutputFormat(str(tmp_path / "p.json"))
    j.writekvs({"a": np.float32(1.5), "b": 2})
    j.close()
    assert json.loads(open(tmp_path / "p.json").read()) == {"a": 1.5, "b": 2}
    df = logger.read_csv(str(tmp_path / "p.csv"))
    assert list(df.columns) == ["a", "b", "c"] and len(df) == 3
    assert logger.read_json(str(tmp_path / "p.json")).iloc[0]["a"] == 1.5


def test_logger_api_mean_dump_levels_and_profile(tmp_path, capsys):
    with logger.scoped_configure(dir=str(tmp_path), format_strs=["stdout", "log", "json", "csv", "tensorboard"]):
        assert logger.get_dir() == str(tmp_path)
        logger.logkv("a", 3)
        logger.logkv_mean("b", -22.5)
        logger.logkv_mean("b", -44.4)
        assert abs(logger.getkvs()["b"] - (-33.45)) < 1e-9   # the reference's demo comment says 33.3
        with logger.ProfileKV("scope"):
            pass
        assert "wait_scope" in logger.getkvs()
        logger.dumpkvs()               # TensorBoard sink is driven by dumpkvs without a crash
        assert dict(logger.getkvs()) == {}
        logger.debug("hidden")
        logger.set_level(logger.DEBUG)
        logger.debug("shown")
        logger.set_level(logger.DISABLED)
        logger.logkv("z", 1)
        logger.dumpkvs()
    out = capsys.readouterr().out
    assert "shown" in out and "hidden" not in out
    assert (tmp_path / "log.txt").exists() and (tmp_path / "progress.csv").exists()
    rows = [json.loads(l) for l in open(tmp_path / "progress.json")]
    assert rows[0]["a"] == 3 and abs(rows[0]["b"] + 33.45) < 1e-6 and len(rows) == 1
    df = logger.read_tb(str(tmp_path / "tb"))
    assert abs(df["b"].iloc[0] + 33.45) < 1e-4 and df["a"].iloc[0] == 3.0

    @logger.profile("fn")
    def f():
        return 5
    assert f() == 5 and logger.getkvs()["wait_fn"] >= 0.0
    logger.getkvs().clear()


def test_tensorboard_format_with_global_step_and_rank_suffix(tmp_path, monkeypatch):
    tb = logger.TensorBoardOutputFormat(str(tmp_path / "run"))
    tb.writekvs({"Global Step": 5, "Loss": 1.5}, global_step=5)   # run_mnist_distributed.py:159
    tb.writekvs({"Loss": 1.25}, global_step=9)
    tb.close()
    df = logger.read_tb(str(tmp_path / "run"))
    assert df.shape == (9, 2) and df["Loss"].iloc[4] == 1.5 and df["Loss"].iloc[8] == 1.25
    assert np.isnan(df["Loss"].iloc[0])
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "3")
    with logger.scoped_configure(dir=str(tmp_path / "mpi")):
        logger.logkv("k", 1)
        logger.dumpkvs()
    assert (tmp_path / "mpi" / "log-rank003.txt").exists()
    with pytest.raises(ValueError):
        logger.make_output_format("nope", str(tmp_path))
