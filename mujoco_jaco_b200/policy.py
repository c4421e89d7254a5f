"""HPC composite-policy forward on the device (BASELINE config 5): the caller of the env-step path.

The reference drives its env with a hierarchical-primitive-composition (HPC) policy of the forked stable-baselines
(`main.py:97-116` primitive wiring, `main.py:159-169` meta policy `weight`, `main.py:252-260` test loop:
`action, subgoal, weight = model.predict_subgoal(obs, deterministic=True)` -> `env.step(action, weight=, subgoal=)`).
The trainer itself is out of scope; what is rebuilt here is the inference path of a saved composite policy:

  * loader for the stable-baselines zip (`models_baseline/policies/*/policy.zip`): JSON `data` (holds the primitive table:
    observation / action index maps, `obs_relativity`, `act_scale`), JSON `parameter_list`, npz `parameters`;
  * the `pi` networks as fp32 PyTorch modules with the saved shapes -- reaching 6 -> 128^3 -> (6 mu, 6 log sigma), grasping
    11 -> 128^3 -> (7, 7), meta policy `weight` 17 -> 256^3 -> (2, 1) (ReLU MLPs, SAC heads `dense` = mu, `dense_1` = log sigma);
  * observation sub-indexing and the `obs_relativity` subtraction per primitive, the weighted composition of the primitive
    Gaussians, tanh squashing and `act_scale`.

[EXT] `stable_baselines.hpc` is not vendored and not pinned: the per-network forward is pinned exactly by the saved
weights (tests/golden/policy_picking.npz holds known-answer vectors evaluated in numpy), the composition rule is restated from
the published HPC / MCP algorithm -- per action dimension j, over the primitives that own that dimension,
    1 / sigma_j^2 = sum_i w_i / sigma_ij^2,        mu_j = sigma_j^2 * sum_i w_i mu_ij / sigma_ij^2,        w = softmax(tanh(meta mu))
(the saved meta network's raw outputs reach +-190: without the SAC squashing a softmax saturates to exactly 0 / 1 and a
dimension owned by one primitive loses its only contributor)
-- and `obs_relativity = {'subtract': {'ref': R, 'tar': T}}` is read as obs[T] - obs[R].  Parity of the composition rule with
the fork is unpinned.
"""
from __future__ import annotations

import io
import json
import re
import zipfile

import numpy as np
import torch

LOG_STD_MIN, LOG_STD_MAX = -20.0, 2.0     # stable-baselines SAC clips log sigma to this range


def load_policy_zip(path):
    """stable-baselines zip -> (data dict, {parameter name: float32 array}) ; only the `model/pi/...` tensors are kept."""
    with zipfile.ZipFile(path) as z:
        data = json.loads(z.read("data"))
        names = json.loads(z.read("parameter_list"))
        npz = np.load(io.BytesIO(z.read("parameters")))
        params = {k: np.asarray(npz[k], np.float32) for k in names if k.startswith("model/pi/")}
    return data, params


def _parse_primitive(info):
    """One entry of data['primitives']: stable-baselines stores it either as a dict or (gym Box objects are not JSON) as the repr
    string of that dict, e.g. "{'obs': (Box(-3.0, 3.0, (23,), float32), [1, 2, ...]), 'act': (Box(...), [0, ...]), 'act_scale': 1,
    'obs_relativity': {'subtract': {'ref': [...], 'tar': [...]}}, ..., 'main_tail': False, ...}"."""
    if isinstance(info, dict):
        if "obs" not in info:
            return None
        rel = (info.get("obs_relativity") or {}).get("subtract")
        return dict(obs_index=list(info["obs"][1]), act_index=list(info["act"][1]), act_scale=info.get("act_scale"),
                    subtract=(list(rel["ref"]), list(rel["tar"])) if rel else None, main_tail=bool(info.get("main_tail")))
    if not isinstance(info, str) or "'obs'" not in info:
        return None
    ints = lambda s: [int(x) for x in s.strip("[]").split(",") if x.strip()]
    obs = re.search(r"'obs': \(.*?\), (\[[0-9, ]*\])\)", info)
    act = re.search(r"'act': \(.*?\), (\[[0-9, ]*\])\)", info)
    if not obs or not act:
        return None
    sc = re.search(r"'act_scale': (None|[0-9.eE+-]+)", info)
    sub = re.search(r"'subtract': \{'ref': (\[[0-9, ]*\]), 'tar': (\[[0-9, ]*\])\}", info)
    return dict(obs_index=ints(obs.group(1)), act_index=ints(act.group(1)),
                act_scale=None if not sc or sc.group(1) == "None" else float(sc.group(1)),
                subtract=(ints(sub.group(1)), ints(sub.group(2))) if sub else None, main_tail="'main_tail': True" in info)


def primitive_table(data):
    """The primitive table of a composite policy: {name: dict(obs_index, act_index, act_scale, subtract=(ref, tar) | None)}."""
    out = {}
    for name, info in data["primitives"].items():
        p = _parse_primitive(info)
        if p is None or p.pop("main_tail"):
            continue
        out[name] = p
    return out


class _Mlp(torch.nn.Module):
    """ReLU MLP with the two SAC heads of stable-baselines (`fc0..fcK`, `dense` = mu, `dense_1` = log sigma)."""

    def __init__(self, sizes, n_mu, n_ls):
        super().__init__()
        self.fc = torch.nn.ModuleList(torch.nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:]))
        self.mu = torch.nn.Linear(sizes[-1], n_mu)
        self.log_std = torch.nn.Linear(sizes[-1], n_ls)

    def forward(self, x):
        for layer in self.fc:
            x = torch.relu(layer(x))
        return self.mu(x), self.log_std(x).clamp(LOG_STD_MIN, LOG_STD_MAX)

    @classmethod
    def from_params(cls, params, prefix):
        ks = sorted(k for k in params if k.startswith(prefix + "/fc") and k.endswith("kernel:0"))
        sizes = [params[ks[0]].shape[0]] + [params[k].shape[1] for k in ks]
        m = cls(sizes, params[prefix + "/dense/kernel:0"].shape[1], params[prefix + "/dense_1/kernel:0"].shape[1])
        with torch.no_grad():
            for i, layer in enumerate(m.fc):
                layer.weight.copy_(torch.from_numpy(params[f"{prefix}/fc{i}/kernel:0"].T)); layer.bias.copy_(torch.from_numpy(params[f"{prefix}/fc{i}/bias:0"]))
            m.mu.weight.copy_(torch.from_numpy(params[prefix + "/dense/kernel:0"].T)); m.mu.bias.copy_(torch.from_numpy(params[prefix + "/dense/bias:0"]))
            m.log_std.weight.copy_(torch.from_numpy(params[prefix + "/dense_1/kernel:0"].T)); m.log_std.bias.copy_(torch.from_numpy(params[prefix + "/dense_1/bias:0"]))
        return m


# the picking composite of main.py:97-116,159-169 (used when no policy file is given: random-initialised weights of the saved shapes)
PICKING_TABLE = {
    "level1_reaching/level0": dict(obs_index=[1, 2, 3, 4, 5, 6, 17, 18, 19, 20, 21, 22], act_index=[0, 1, 2, 3, 4, 5], act_scale=1,
                                   subtract=([17, 18, 19, 20, 21, 22], [1, 2, 3, 4, 5, 6])),
    "level1_grasping/level0": dict(obs_index=list(range(11)), act_index=list(range(7)), act_scale=1, subtract=None),
    "level1_picking/weight": dict(obs_index=list(range(11)) + list(range(17, 23)), act_index=[0, 1], act_scale=None, subtract=None),
}
PICKING_SHAPES = {"level1_reaching/level0": ([6, 128, 128, 128], 6, 6), "level1_grasping/level0": ([11, 128, 128, 128], 7, 7),
                  "level1_picking/weight": ([17, 256, 256, 256], 2, 1)}


class CompositePolicy(torch.nn.Module):
    """Composite HPC policy: primitive Gaussians blended by the meta policy's weights.  `predict_subgoal(obs, deterministic)`
    mirrors the call of main.py:259 and returns (action [B, 7], subgoal dict, weight [B, n_primitives])."""

    def __init__(self, table, nets, n_act=7):
        super().__init__()
        self.table = table
        self.names = [k for k in table if not k.endswith("/weight")]
        self.weight_name = next(k for k in table if k.endswith("/weight"))
        self.nets = torch.nn.ModuleDict({k.replace("/", "__"): v for k, v in nets.items()})
        self.n_act = n_act

    @classmethod
    def from_zip(cls, path):
        data, params = load_policy_zip(path)
        table = primitive_table(data)
        return cls(table, {k: _Mlp.from_params(params, "model/pi/" + k) for k in table})

    @classmethod
    def from_arrays(cls, table, params):
        return cls(table, {k: _Mlp.from_params(params, "model/pi/" + k) for k in table})

    @classmethod
    def random_init(cls, seed=0):
        """Picking composite with the saved architecture and random weights (bench.py config 5: no checkpoints on the GPU box)."""
        g = torch.Generator().manual_seed(seed)
        nets = {}
        for k, (sizes, n_mu, n_ls) in PICKING_SHAPES.items():
            m = _Mlp(sizes, n_mu, n_ls)
            with torch.no_grad():
                for p in m.parameters():
                    p.copy_(torch.randn(p.shape, generator=g) * (0.1 if p.dim() > 1 else 0.01))
            nets[k] = m
        return cls(dict(PICKING_TABLE), nets)

    def primitive_input(self, name, obs):
        info = self.table[name]
        if info["subtract"] is not None:
            ref, tar = info["subtract"]
            return obs[:, tar] - obs[:, ref]
        return obs[:, info["obs_index"]]

    def forward(self, obs, deterministic=True, generator=None):
        B, dev = obs.shape[0], obs.device
        w_mu, _ = self.nets[self.weight_name.replace("/", "__")](self.primitive_input(self.weight_name, obs))
        weight = torch.softmax(torch.tanh(w_mu), dim=1)      # SAC deterministic output tanh(mu) in (-1, 1), normalised over the primitives
        prec = torch.zeros(B, self.n_act, device=dev)      # sum_i w_i / sigma_ij^2
        pmu = torch.zeros(B, self.n_act, device=dev)       # sum_i w_i mu_ij / sigma_ij^2
        scale = torch.ones(self.n_act, device=dev)
        for i, name in enumerate(self.names):
            mu, log_std = self.nets[name.replace("/", "__")](self.primitive_input(name, obs))
            idx = self.table[name]["act_index"]
            p = weight[:, i:i + 1] * torch.exp(-2.0 * log_std)
            prec[:, idx] += p
            pmu[:, idx] += p * mu
            if self.table[name]["act_scale"] is not None:
                scale[idx] = float(self.table[name]["act_scale"])
        prec = prec.clamp_min(1e-12)
        mu_c, std_c = pmu / prec, prec.rsqrt()
        pre = mu_c if deterministic else mu_c + std_c * torch.randn(mu_c.shape, device=dev, generator=generator)
        return torch.tanh(pre) * scale, weight, mu_c, std_c

    @torch.no_grad()
    def predict_subgoal(self, obs, deterministic=True):
        """obs: [26] / [B, 26] numpy array or torch tensor.  Returns (action, subgoal, weight) like HPC.predict_subgoal: subgoal is the
        dict the reference passes through to env.step (empty for the rule-based subgoal of every main.py entry point)."""
        is_np = isinstance(obs, np.ndarray)
        x = torch.as_tensor(obs, dtype=torch.float32, device=next(self.parameters()).device)
        single = x.dim() == 1
        if single:
            x = x[None]
        action, weight, _, _ = self.forward(x, deterministic)
        if single:
            action, weight = action[0], weight[0]
        if is_np:
            return action.cpu().numpy(), {}, weight.cpu().numpy()
        return action, {}, weight
