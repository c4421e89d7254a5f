"""Generates tests/golden/policy_picking.npz from the reference's saved composite policy
(/root/reference/models_baseline/policies/picking/policy.zip): the `model/pi/...` tensors of the three networks, the primitive
table of its `data` entry (as JSON), 64 seeded observations and the known answers -- every network's (mu, log sigma) and the
composite (action, weight) -- evaluated here in plain float64 numpy, independently of mujoco_jaco_b200/policy.py.
Run in the build container (the reference tree does not exist on the GPU box): python tests/golden/make_policy_golden.py"""
import io, json, os, sys, zipfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/models_baseline/policies/picking/policy.zip"


def mlp(p, prefix, x):
    k = 0
    while f"{prefix}/fc{k}/kernel:0" in p:
        x = np.maximum(x @ p[f"{prefix}/fc{k}/kernel:0"].astype(np.float64) + p[f"{prefix}/fc{k}/bias:0"], 0.0)
        k += 1
    mu = x @ p[prefix + "/dense/kernel:0"].astype(np.float64) + p[prefix + "/dense/bias:0"]
    ls = np.clip(x @ p[prefix + "/dense_1/kernel:0"].astype(np.float64) + p[prefix + "/dense_1/bias:0"], -20.0, 2.0)
    return mu, ls


def main():
    z = zipfile.ZipFile(SRC)
    data = json.loads(z.read("data"))
    names = [k for k in json.loads(z.read("parameter_list")) if k.startswith("model/pi/")]
    npz = np.load(io.BytesIO(z.read("parameters")))
    p = {k: np.asarray(npz[k], np.float32) for k in names}
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from mujoco_jaco_b200.policy import primitive_table          # only the parser of the `data` entry; the arithmetic below is independent
    table = {k: dict(v, subtract=list(v["subtract"]) if v["subtract"] else None) for k, v in primitive_table(data).items()}
    rng = np.random.default_rng(2024)
    obs = rng.uniform(-1, 1, (64, 26)).astype(np.float32)
    obs[:, 0] = rng.integers(0, 4, 64)
    out = {}
    x64 = obs.astype(np.float64)
    inp = lambda n: (x64[:, table[n]["subtract"][1]] - x64[:, table[n]["subtract"][0]]) if table[n]["subtract"] else x64[:, table[n]["obs_index"]]
    wname = next(k for k in table if k.endswith("/weight"))
    wmu, wls = mlp(p, "model/pi/" + wname, inp(wname))
    tw = np.tanh(wmu)
    w = np.exp(tw - tw.max(1, keepdims=True)); w /= w.sum(1, keepdims=True)
    prec, pmu = np.zeros((64, 7)), np.zeros((64, 7))
    for i, n in enumerate(k for k in table if not k.endswith("/weight")):
        mu, ls = mlp(p, "model/pi/" + n, inp(n))
        out[f"ans_mu_{i}"], out[f"ans_logstd_{i}"] = mu, ls
        idx = table[n]["act_index"]
        pr = w[:, i:i + 1] * np.exp(-2 * ls)
        prec[:, idx] += pr; pmu[:, idx] += pr * mu
    out.update(ans_weight_mu=wmu, ans_weight=w, ans_action=np.tanh(pmu / prec), ans_mu_c=pmu / prec, ans_std_c=prec ** -0.5)
    np.savez_compressed(os.path.join(HERE, "policy_picking.npz"), obs=obs, table=json.dumps(table), **{"p:" + k: v for k, v in p.items()}, **out)
    print("wrote policy_picking.npz:", {k: v.shape for k, v in p.items() if "fc0/kernel" in k}, "weights (first rows)", w[:3])


if __name__ == "__main__":
    main()
