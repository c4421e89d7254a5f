"""HPC composite-policy forward (BASELINE config 5): mujoco_jaco_b200/policy.py against known-answer vectors
evaluated in float64 numpy from the reference's saved weights (tests/golden/policy_picking.npz, made by make_policy_golden.py)."""
import io
import json
import os
import zipfile

import numpy as np
import pytest

torch = pytest.importorskip("torch")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "policy_picking.npz")


def _load():
    from mujoco_jaco_b200.policy import CompositePolicy
    Z = np.load(GOLD)
    table = json.loads(str(Z["table"]))
    for v in table.values():
        v["subtract"] = tuple(v["subtract"]) if v["subtract"] else None
    params = {k[2:]: Z[k] for k in Z.files if k.startswith("p:")}
    return Z, table, params, CompositePolicy.from_arrays(table, params)


def _check(pol, Z, dev):
    pol = pol.to(dev).requires_grad_(False)
    obs = torch.tensor(Z["obs"], device=dev)
    action, weight, mu_c, std_c = pol(obs, deterministic=True)
    assert np.allclose(weight.cpu().numpy(), Z["ans_weight"], rtol=1e-5, atol=1e-6)
    assert np.allclose(mu_c.detach().cpu().numpy(), Z["ans_mu_c"], rtol=1e-4, atol=2e-5)
    assert np.allclose(action.detach().cpu().numpy(), Z["ans_action"], rtol=1e-4, atol=2e-5)
    for i, name in enumerate(pol.names):
        mu, ls = pol.nets[name.replace("/", "__")](pol.primitive_input(name, obs))
        assert np.allclose(mu.detach().cpu().numpy(), Z[f"ans_mu_{i}"], rtol=1e-4, atol=2e-5)
        assert np.allclose(ls.detach().cpu().numpy(), Z[f"ans_logstd_{i}"], rtol=1e-4, atol=2e-5)
    a, sg, w = pol.predict_subgoal(Z["obs"][0], deterministic=True)            # the call of main.py:259
    assert a.shape == (7,) and w.shape == (2,) and sg == {} and np.allclose(a, Z["ans_action"][0], atol=2e-5)
    assert (np.abs(a) <= 1).all() and abs(w.sum() - 1) < 1e-6


def test_composite_forward_matches_numpy_known_answers_cpu():
    Z, table, params, pol = _load()
    assert [tuple(params[f"model/pi/{k}/fc0/kernel:0"].shape) for k in table] == [(6, 128), (11, 128), (17, 256)]   # SURVEY 8d C5 shapes
    assert table["level1_reaching/level0"]["subtract"] == ([17, 18, 19, 20, 21, 22], [1, 2, 3, 4, 5, 6])          # main.py:100-102
    _check(pol, Z, "cpu")


def test_policy_zip_loader_round_trip(tmp_path):
    """A stable-baselines style zip (JSON data + parameter_list + npz parameters) written from the golden arrays loads into the
    same policy (the reference's own policy.zip files are not on the GPU box)."""
    from mujoco_jaco_b200.policy import CompositePolicy
    Z, table, params, pol = _load()
    prim = {k: {"obs": [None, v["obs_index"]], "act": [None, v["act_index"]], "act_scale": v["act_scale"],
                "obs_relativity": {"subtract": {"ref": v["subtract"][0], "tar": v["subtract"][1]}} if v["subtract"] else {}} for k, v in table.items()}
    prim["continue"] = {"obs": [None, list(range(23))], "act": [None, list(range(7))], "main_tail": True}
    buf = io.BytesIO(); np.savez(buf, **params)
    path = tmp_path / "policy.zip"
    with zipfile.ZipFile(path, "w") as z:
        z.writestr("data", json.dumps({"primitives": prim, "composite_primitive_name": "picking"}))
        z.writestr("parameter_list", json.dumps(list(params)))
        z.writestr("parameters", buf.getvalue())
    _check(CompositePolicy.from_zip(str(path)), Z, "cpu")


def test_random_init_policy_has_the_saved_architecture():
    from mujoco_jaco_b200.policy import CompositePolicy
    pol = CompositePolicy.random_init(seed=3).requires_grad_(False)
    a, w, mu, std = pol(torch.zeros(5, 26))
    assert a.shape == (5, 7) and w.shape == (5, 2) and torch.isfinite(a).all() and (std > 0).all()
    n = sum(p.numel() for p in pol.parameters())
    assert n == sum(v.size for k, v in _load()[2].items())


@pytest.mark.gpu
def test_composite_forward_on_device_and_policy_in_the_loop():
    """The forward on cuda:0 against the numpy answers, then 3 steps of policy -> jaco_step with no host round trip."""
    from mujoco_jaco_b200.env import JacoVecEnv
    Z, table, params, pol = _load()
    _check(pol, Z, "cuda")
    n = 1024
    env = JacoVecEnv(n, task="picking", seed=1, auto_reset=True)
    obs = env.reset()
    for _ in range(3):
        with torch.no_grad():
            action, weight, _, _ = pol(obs, deterministic=True)
        obs, rew, done, succ = env.step(action)
        assert torch.isfinite(obs).all() and torch.isfinite(rew).all() and action.abs().max() <= 1
