"""`.pth` compatibility: the archive written by ppo_cpp_b200.pth_compat carries the reference agent's
parameter names (src/ppo_continuous_action.cpp:120-137), opens with torch.jit.load (the reader for `torch::save(agent)`
files, :173-180), evaluates to the same value / mean as the torch oracle agent, and round-trips to the flat vector
bit-exactly; the engine's own checkpoint header is parsed without the GPU."""
import struct
import zipfile

import numpy as np
import pytest
import torch

from oracle import oracle_nn
from ppo_cpp_b200 import pth_compat


@pytest.mark.parametrize("obs,act", [(11, 3), (376, 17)])
def test_export_roundtrip_and_names(tmp_path, obs, act):
    rng = np.random.default_rng(obs)
    flat = rng.normal(size=pth_compat.num_params(obs, act)).astype(np.float32)
    path = tmp_path / "model_final.pth"
    pth_compat.save_pth(path, flat, obs, act)
    assert zipfile.is_zipfile(path)                       # torch::save archives are zip files with data.pkl inside
    assert any(n.endswith("data.pkl") for n in zipfile.ZipFile(path).namelist())
    mod = torch.jit.load(str(path))
    names = [n for n, _ in mod.named_parameters()]
    assert sorted(names) == sorted(pth_compat.NAMES)
    assert np.array_equal(pth_compat.load_pth(path), flat)
    # same numbers as the oracle agent built from the same flat vector (layout = AgentImpl's registration order)
    agent = oracle_nn.Agent(obs, act)
    agent.load_flat(flat)
    x = torch.from_numpy(rng.normal(size=(5, obs)).astype(np.float32))
    sd = dict(mod.named_parameters())
    h = torch.tanh(x @ sd["critic.0.weight"].T + sd["critic.0.bias"])
    h = torch.tanh(h @ sd["critic.2.weight"].T + sd["critic.2.bias"])
    v = h @ sd["critic.4.weight"].T + sd["critic.4.bias"]
    assert torch.allclose(v, agent.get_value(x), atol=1e-6)


def test_plain_state_dict_is_accepted(tmp_path):
    obs, act = 17, 6
    flat = np.arange(pth_compat.num_params(obs, act), dtype=np.float32)
    torch.save(pth_compat.state_dict_from_flat(flat, obs, act), tmp_path / "sd.pth")
    assert np.array_equal(pth_compat.load_pth(tmp_path / "sd.pth"), flat)


def test_engine_checkpoint_header(tmp_path):
    obs, act, N = 11, 3, 4
    P = pth_compat.num_params(obs, act)
    body = np.arange(3 * P, dtype=np.float32)
    p = tmp_path / "model_final.bin"
    with open(p, "wb") as f:
        f.write(struct.pack("<8siiiif3i", b"PB200CK\0", 1, P, N, obs, 7.0, 0, 0, 0))
        body.tofile(f)
        np.zeros(2 * N * obs + 5 * N, np.float32).tofile(f)     # normaliser statistics follow, ignored here
    ck = pth_compat.read_checkpoint(p)
    assert ck["P"] == P and ck["O"] == obs and ck["step"] == 7.0
    assert np.array_equal(ck["params"], body[:P]) and np.array_equal(ck["adam_v"], body[2 * P:])
    assert pth_compat.main(["export", str(p), str(obs), str(act), str(tmp_path / "m.pth")]) == 0
    assert np.array_equal(pth_compat.load_pth(tmp_path / "m.pth"), body[:P])
    with pytest.raises(ValueError):
        (tmp_path / "bad.bin").write_bytes(b"x" * 64)
        pth_compat.read_checkpoint(tmp_path / "bad.bin")
