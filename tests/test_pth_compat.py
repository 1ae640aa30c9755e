"""`.pth` compatibility: the archive written by ppo_cpp_b200.pth_compat carries the reference agent's
parameter names (src/ppo_continuous_action.cpp:120-137), opens with torch.jit.load (the reader for `torch::save(agent)`
files, :173-180), evaluates to the same value / mean as the torch oracle agent, and round-trips to the flat vector
bit-exactly; the engine's own checkpoint header is parsed without the GPU."""
import os
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


PTH_WRITER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "pth_writer")


@pytest.mark.skipif(not os.path.exists(PTH_WRITER), reason="build/pth_writer not built (needs a libtorch at build time)")
@pytest.mark.parametrize("kind,obs,act,P", [(0, 376, 17, 57763), (0, 11, 3, 10119), (1, 17, 6, 146189)])
def test_pth_writer_round_trip_through_libtorch(tmp_path, kind, obs, act, P):
    """host/pth_writer.cpp: torch::save(agent) + torch::save(optimizer) from the CLIs' raw dump, then - as a reference-side tool
    would - torch::load into freshly constructed AgentImpl / optim::Adam objects: parameters, exp_avg, exp_avg_sq and step come
    back bit-exact.  The parameter counts are SURVEY.md 8a row A14 / A15's."""
    import struct
    import subprocess
    rng = np.random.default_rng(kind)
    raw, model, opt = (str(tmp_path / n) for n in ("d.raw", "model.pth", "optimizer.pth"))
    params = rng.normal(size=P).astype(np.float32)
    with open(raw, "wb") as f:
        f.write(struct.pack("<iiiii", 0x50425257, kind, obs, act, P) + struct.pack("<q", 320) + struct.pack("<ffff", 3e-4, 1e-5, 0.4, -0.4))
        f.write(params.tobytes() + rng.normal(size=P).astype(np.float32).tobytes() + np.abs(rng.normal(size=P)).astype(np.float32).tobytes())
        if kind == 1:
            f.write(rng.normal(size=obs).astype(np.float32).tobytes() + (np.abs(rng.normal(size=obs)) + 0.1).astype(np.float32).tobytes())
    for mode in ("write", "verify"):
        r = subprocess.run([PTH_WRITER, mode, raw, model, opt], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, (mode, r.stdout, r.stderr[-1000:])
    assert "params_max_err 0.000e+00 adam_max_err 0.000e+00" in r.stdout and "step_ok" in r.stdout
    m = torch.jit.load(model)
    named = dict(m.named_parameters())
    if kind == 0:
        assert set(named) == set(pth_compat.NAMES)
        flat = np.concatenate([named[n].detach().numpy().ravel() for n in pth_compat.NAMES])
        assert np.array_equal(flat, params)
    else:
        assert {"mean_", "std_", "action_space_high", "action_space_low", "critic.0.weight", "actor_mean.4.bias", "dist_alpha.0.weight",
                "dist_beta.0.bias"} <= set(named)
