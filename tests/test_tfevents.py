"""TensorBoard event file written by include/tfevents_writer.h (reference:
libs/tensorboard_logger/src/tensorboard_logger.cc:305-327 record framing, src/crc.cc:235-238 CRC masking,
tensorboard_logger.cc:164-184 text summaries).  The file is checked twice: record framing and both masked CRC32C fields
with an independent Python restatement, and the decoded events with TensorBoard's own loader (the consumer the
reference's files are made for)."""
import os
import struct
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _crc32c(data):
    table = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        table.append(c)
    crc = 0xFFFFFFFF
    for b in data:
        crc = table[(crc ^ b) & 0xFF] ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def _masked(data):
    crc = _crc32c(data)
    return (((crc >> 15) | (crc << 17)) + 0xA282EAD8) & 0xFFFFFFFF


@pytest.fixture(scope="module")
def event_file(tmp_path_factory):
    exe = os.path.join(ROOT, "build", "tfevents_selftest")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++20", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "host", "tfevents_selftest.cpp")])
    path = str(tmp_path_factory.mktemp("tb") / "tfevents_logs.pb")   # the reference's file name (:281)
    subprocess.check_call([exe, path])
    return path


def test_crc32c_known_answers():
    assert _crc32c(b"123456789") == 0xE3069283          # the standard CRC-32C check value
    assert _crc32c(b"\x00" * 32) == 0x8A9136AA          # RFC 3720 B.4


def test_record_framing_and_masked_crcs(event_file):
    raw = open(event_file, "rb").read()
    off, n = 0, 0
    while off < len(raw):
        (length,) = struct.unpack_from("<Q", raw, off)
        (len_crc,) = struct.unpack_from("<I", raw, off + 8)
        body = raw[off + 12: off + 12 + length]
        (data_crc,) = struct.unpack_from("<I", raw, off + 12 + length)
        assert len_crc == _masked(raw[off: off + 8]) and data_crc == _masked(body)
        off += 12 + length + 4
        n += 1
    assert off == len(raw) and n == 13                  # version record + 1 text + 11 scalars


def test_tensorboard_reads_tags_steps_and_values(event_file):
    loader = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    evs = list(loader.LegacyEventFileLoader(event_file).Load())
    assert evs[0].file_version == "brain.Event:2"
    text = evs[1].summary.value[0]
    assert text.tag == "hyperparameters" and text.metadata.plugin_data.plugin_name == "text"
    assert text.tensor.string_val[0] == b"env_id: Humanoid-v4\nnum_envs: 8192\n" and text.tensor.dtype == 7
    scal = [(e.step, e.summary.value[0].tag, e.summary.value[0].simple_value) for e in evs[2:]]
    assert scal[0] == (262144, "charts/SPS", 1500001.0) and scal[1] == (262144, "losses/value_loss", 0.25)
    assert scal[-1] == (3000000000, "charts/episodic_return", -12.5)
    assert all(e.wall_time > 1.6e9 for e in evs)


def test_file_name_must_contain_tfevents(event_file, tmp_path):
    exe = os.path.join(ROOT, "build", "tfevents_selftest")
    r = subprocess.run([exe, str(tmp_path / "events.pb")], capture_output=True, text=True)
    assert r.returncode == 1 and "tfevents" in r.stderr
